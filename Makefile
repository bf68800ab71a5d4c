# Top-level build: libfxgpu.so (CUDA, sm_100a only), the fxtract host CLI, and the oracle.
ROOT := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))
BLD := $(ROOT)fxtract_b200/_build
CSRC := $(ROOT)fxtract_b200/csrc
HOST := $(ROOT)fxtract_b200/host
NVCC ?= nvcc
CXX ?= g++
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -Wall
CU := $(CSRC)/fxg_scan.cu $(CSRC)/fxg_scan_kmer.cu $(CSRC)/fxg_fields.cu $(CSRC)/fxg_synth.cu $(CSRC)/fxg_api.cu
CUOBJ := $(patsubst $(CSRC)/%.cu,$(BLD)/%.o,$(CU))
CUHDR := $(wildcard $(CSRC)/*.cuh) $(ROOT)include/fxgpu.h
HOSTSRC := $(wildcard $(HOST)/*.cpp)

.PHONY: all lib cli tools oracle clean fxtract_test ptxas
all: lib cli tools

# stand-alone generator of the synthetic workload (no CUDA): bench.py's reference arm and the verification legs
tools: $(BLD)/fxsynth
$(BLD)/fxsynth: $(ROOT)fxtract_b200/tools/fxsynth.cpp $(CSRC)/fxg_synth_core.h
	@mkdir -p $(BLD)
	$(CXX) -O2 -std=c++17 -Wall -pthread -o $@ $<

lib: $(BLD)/libfxgpu.so
$(BLD)/%.o: $(CSRC)/%.cu $(CUHDR)
	@mkdir -p $(BLD)
	$(NVCC) $(NVFLAGS) -c -o $@ $<
$(BLD)/libfxgpu.so: $(CUOBJ)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CUOBJ)

cli: $(BLD)/fxtract
$(BLD)/fxtract: $(HOSTSRC) $(wildcard $(HOST)/*.h) $(BLD)/libfxgpu.so
	@if [ -n "$(HOSTSRC)" ]; then \
	  $(CXX) -O2 -std=c++17 -Wall -I$(ROOT)include -o $@ $(HOSTSRC) -L$(BLD) -lfxgpu -lz -Wl,-rpath,'$$ORIGIN'; \
	else echo "host CLI sources not present yet"; fi

ptxas:
	$(NVCC) $(NVFLAGS) -Xptxas -v -cubin -o /dev/null $(CSRC)/fxg_scan.cu

oracle:
	$(MAKE) -C $(ROOT)oracle port ref

# the reference's own acceptance entry point (Makefile:56-57): needs a GPU
fxtract_test: all
	cd $(ROOT) && python -m pytest tests/test_cli_golden.py -q -m gpu

clean:
	rm -rf $(BLD)
