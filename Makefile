# Builds libodex.so (the C-ABI CUDA library) in-tree for sm_100a.  `python -c "import __graft_entry__ as g; g.build()"`
# drives this; the .so is git-ignored but travels to the GPU box with gpurun.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC -Xcompiler -Wall -Xptxas -v
CSRC := ode_rl_b200/csrc
OBJS := $(CSRC)/build/api.o $(CSRC)/build/conv_simt.o $(CSRC)/build/conv_tc.o $(CSRC)/build/gru.o $(CSRC)/build/rk.o $(CSRC)/build/wgrad_tc.o $(CSRC)/build/conv_tc2.o $(CSRC)/build/convnet_tc.o $(CSRC)/build/convnet_f16.o $(CSRC)/build/conv16_f16.o $(CSRC)/build/wgrad_f16.o $(CSRC)/build/vid.o $(CSRC)/build/decode_tail.o $(CSRC)/build/upsample.o $(CSRC)/build/encode_stem.o
LIB := ode_rl_b200/libodex.so

all: $(LIB)

$(CSRC)/build/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/odex.h
	@mkdir -p $(CSRC)/build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(CSRC)/build/$*.ptxas.log || (cat $(CSRC)/build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -shared $(ARCH) -o $@ $(OBJS) -cudart static

clean:
	rm -rf $(CSRC)/build $(LIB)
.PHONY: all clean
