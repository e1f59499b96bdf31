# Builds libcor_b200.so (the C-ABI product library) for sm_100a.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
CSRC := cor_asv_ann_b200/csrc
SRCS := $(CSRC)/api.cu $(CSRC)/kernels.cu $(CSRC)/attention.cu $(CSRC)/gemm_ffma.cu $(CSRC)/gemm_tc.cu $(CSRC)/beam.cu $(CSRC)/alignpath.cu $(CSRC)/recur.cu $(CSRC)/decpage.cu $(CSRC)/train.cu $(CSRC)/noalloc.cu
OBJS := $(SRCS:.cu=.o)
HDRS := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/cor_b200.h
LIB  := cor_asv_ann_b200/libcor_b200.so
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr

all: $(LIB)

%.o: %.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart -lcublas -lpthread

clean:
	rm -f $(OBJS) $(LIB)
.PHONY: all clean
