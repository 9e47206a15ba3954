"""Drop-in for Models/GPT2/config.py (Models/GPT2/config.py:6-50): same keys and sizes."""
_SIZES = {"124M": (768, 12, 12), "355M": (1024, 16, 24), "774M": (1280, 20, 36), "1.5B": (1600, 25, 48)}


def build_gpt_config(emb_dim, n_heads, n_layers):
    return {"vocab_size": 50257, "context_length": 1024, "emb_dim": emb_dim, "n_heads": n_heads,
            "n_layers": n_layers, "drop_rate": 0.1, "qkv_bias": False, "eos_id": 50256, "eos_text": "<|endoftext|>"}


available_configs = {name: build_gpt_config(*dims) for name, dims in _SIZES.items()}


def get_config_gpt2(num_params):
    num_params = str(num_params)
    if num_params not in available_configs:
        raise ValueError(f"GPT-2 config for model '{num_params}' not found. "
                         f"Available options: {list(available_configs.keys())}")
    return available_configs[num_params]
