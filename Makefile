# Builds libisr.so (sm_100a only) in-tree and the CPU oracle (test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH = -gencode arch=compute_100a,code=sm_100a
PKG = isosurfacesuperresolution_b200
CSRC = $(PKG)/csrc
NVFLAGS = $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v
LIB = $(PKG)/libisr.so

all: $(LIB) oracle
$(CSRC)/isr_api.o: $(CSRC)/isr_api.cu $(CSRC)/isr_conv.cuh $(CSRC)/isr_conv_final.cuh $(CSRC)/isr_conv_final2.cuh $(CSRC)/isr_conv_pair.cuh $(CSRC)/isr_ptx.cuh $(CSRC)/isr_internal.h include/isr.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(CSRC)/isr_pointwise.o: $(CSRC)/isr_pointwise.cu $(CSRC)/isr_internal.h
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(CSRC)/isr_ingest.o: $(CSRC)/isr_ingest.cu $(CSRC)/isr_internal.h
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(CSRC)/isr_up2.o: $(CSRC)/isr_up2.cu $(CSRC)/isr_internal.h
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(CSRC)/isr_ssao.o: $(CSRC)/isr_ssao.cu $(CSRC)/isr_internal.h
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(CSRC)/isr_raycast.o: $(CSRC)/isr_raycast.cu $(CSRC)/isr_internal.h include/isr.h
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(LIB): $(CSRC)/isr_raycast.o $(CSRC)/isr_api.o $(CSRC)/isr_pointwise.o $(CSRC)/isr_ingest.o $(CSRC)/isr_ssao.o $(CSRC)/isr_up2.o
	$(NVCC) $(ARCH) -shared -o $@ $^ -cudart static
oracle:
	$(MAKE) -C oracle -s
clean:
	rm -f $(CSRC)/*.o $(LIB)
.PHONY: all oracle clean
