"""Drop-in for Models/Llama/config.py: the same configuration dictionaries and loader.

get_config_llama keeps the reference's behaviour (context forced to 1024 with theta rescaled,
Models/Llama/config.py:114-124) but works on a copy instead of mutating the module-level dict,
and takes an optional target_context_length so cfg #3 (2048) / #5 (4096) can be requested
(SURVEY.md R4)."""
import copy

import torch

from .common_components import rescale_theta

_LLAMA3_FREQ_8192 = {"low_freq_factor": 1.0, "high_freq_factor": 4.0, "original_context_length": 8192}

LLAMA2_CONFIG_7B = {"vocab_size": 32_000, "context_length": 4096, "emb_dim": 4096, "n_heads": 32, "n_layers": 32,
                    "hidden_dim": 11_008, "dtype": torch.bfloat16}

LLAMA3_CONFIG_8B = {"vocab_size": 128_256, "context_length": 8192, "emb_dim": 4096, "n_heads": 32, "n_layers": 32,
                    "hidden_dim": 14_336, "n_kv_groups": 8, "rope_base": 500_000.0, "rope_freq": None,
                    "dtype": torch.bfloat16, "eos_id": 128_001, "eos_text": "<|end_of_text|>"}

LLAMA31_CONFIG_8B = {"vocab_size": 128_256, "context_length": 131_072, "emb_dim": 4096, "n_heads": 32, "n_layers": 32,
                     "hidden_dim": 14_336, "n_kv_groups": 8, "rope_base": 500_000.0, "dtype": torch.bfloat16,
                     "rope_freq": dict(_LLAMA3_FREQ_8192, factor=8.0), "eos_id": 128_001,
                     "eos_text": "<|end_of_text|>"}

LLAMA32_CONFIG_1B = {"vocab_size": 128_256, "context_length": 131_072, "emb_dim": 2048, "n_heads": 32, "n_layers": 16,
                     "hidden_dim": 8192, "n_kv_groups": 8, "rope_base": 500_000.0, "dtype": torch.bfloat16,
                     "rope_freq": dict(_LLAMA3_FREQ_8192, factor=32.0), "eos_id": 128_001,
                     "eos_text": "<|end_of_text|>"}

available_configs_llama2 = {"7B": LLAMA2_CONFIG_7B}
available_configs_llama3 = {"8B": LLAMA3_CONFIG_8B}
available_configs_llama3_1 = {"8B": LLAMA31_CONFIG_8B}
available_configs_llama3_2 = {"1B": LLAMA32_CONFIG_1B}


def get_config_llama(num_params, model_name, target_context_length=1024):
    num_params = str(num_params)
    table = globals().get(f"available_configs_{model_name}", None)
    assert table is not None, f"No configuration found for model '{model_name}'"
    assert num_params in table, f"A {model_name} model with {num_params} parameters does not exist."
    config = copy.deepcopy(table[num_params])
    old = config["context_length"]
    if old != target_context_length:
        config["context_length"] = target_context_length
        config["rope_base"] = rescale_theta(config["rope_base"], old, target_context_length)
    return config
