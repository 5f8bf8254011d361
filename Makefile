# Build libmpb.so (sm_100a CUDA + C-ABI) and the CPU oracle (test infrastructure).
NVCC ?= nvcc
NVCCFLAGS = -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-Wall -shared
LIB = mplib_b200/libmpb.so
SRC = mplib_b200/csrc/mpb.cu
HDR = $(wildcard mplib_b200/csrc/*.cuh) include/mpb.h

all: $(LIB) oracle

$(LIB): $(SRC) $(HDR)
	mkdir -p build
	$(NVCC) $(NVCCFLAGS) -Xptxas -v -o $@ $(SRC) 2> build/ptxas.log || (cat build/ptxas.log; false)

oracle: oracle/_build/liboracle.so
oracle/_build/liboracle.so: oracle/oracle.c include/mpb.h
	mkdir -p oracle/_build
	gcc -O3 -fopenmp -fPIC -shared -Wall -Wextra -o $@ oracle/oracle.c -lm

clean:
	rm -f $(LIB) oracle/_build/liboracle.so

.PHONY: all oracle clean
