# Builds libxcorr_b200.so (CUDA kernels + C-ABI, sm_100a only) and the oracle-side helpers.
#   make            -> pydsproutines_b200/libxcorr_b200.so
#   make clean
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC \
             --expt-relaxed-constexpr -Xptxas -v $(EXTRA)
# EXTRA=-DXC_EXPERIMENTS keeps the earlier layouts of the N = 8192 search selectable (XC_LAYOUT=b2|dual|staged|pp)
# for A/B runs; the shipped library is built without it.
CSRC      := pydsproutines_b200/csrc
BUILD     := build
SRCS      := $(wildcard $(CSRC)/*.cu)
OBJS      := $(patsubst $(CSRC)/%.cu,$(BUILD)/%.o,$(SRCS))
HDRS      := $(wildcard $(CSRC)/*.cuh) include/xcorr_b200.h
LIB       := pydsproutines_b200/libxcorr_b200.so

all: $(LIB)

$(BUILD)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(BUILD)/$*.ptxas.log || (cat $(BUILD)/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static -ldl

clean:
	rm -rf $(BUILD) $(LIB)

.PHONY: all clean
