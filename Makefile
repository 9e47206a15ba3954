# Builds the C-ABI library (sm_100a only) and the oracle's C helpers.
NVCC ?= nvcc
PKG := building_llm_from_scratch_b200
CSRC := $(PKG)/csrc
OBJDIR := build/obj
SRCS := $(wildcard $(CSRC)/*.cu)
OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(SRCS))
NVCCFLAGS := -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo \
             -Xcompiler -fPIC -Xcompiler -fvisibility=default --expt-relaxed-constexpr \
             -Xptxas -v -Iinclude

LIB := $(PKG)/libb200llm.so

all: $(LIB)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(CSRC)/common.cuh include/b200llm.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $(OBJS) -lcudart

clean:
	rm -rf build $(LIB)

.PHONY: all clean
