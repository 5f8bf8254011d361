# Build libmpb.so (sm_100a CUDA + C-ABI) and the CPU oracle (test infrastructure).
NVCC ?= nvcc
NVCCFLAGS = -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-Wall -shared
LIB = mplib_b200/libmpb.so
SRC = mplib_b200/csrc/mpb.cu
HDR = $(wildcard mplib_b200/csrc/*.cuh) include/mpb.h

all: $(LIB) native oracle

$(LIB): $(SRC) $(HDR)
	mkdir -p build
	$(NVCC) $(NVCCFLAGS) -Xptxas -v -o $@ $(SRC) 2> build/ptxas.log || (cat build/ptxas.log; false)

# pybind11 host module over the C-ABI (per-call path): g++ only, Python.h + pybind11 headers from the interpreter on PATH
PYEXT := $(shell python -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
PYINC := $(shell python -c "import sysconfig, pybind11; print('-I' + sysconfig.get_paths()['include'] + ' -I' + pybind11.get_include())")
NATIVE = mplib_b200/_native$(PYEXT)
native: $(NATIVE)
$(NATIVE): mplib_b200/csrc/pymp_native.cpp include/mpb.h $(LIB)
	g++ -O2 -std=c++17 -fPIC -shared -fvisibility=hidden -Wall $(PYINC) -o $@ mplib_b200/csrc/pymp_native.cpp -Lmplib_b200 -lmpb -Wl,-rpath,'$$ORIGIN'

oracle: oracle/_build/liboracle.so
oracle/_build/liboracle.so: oracle/oracle.c include/mpb.h
	mkdir -p oracle/_build
	gcc -O3 -fopenmp -fPIC -shared -Wall -Wextra -o $@ oracle/oracle.c -lm

clean:
	rm -f $(LIB) $(NATIVE) oracle/_build/liboracle.so

.PHONY: all native oracle clean
