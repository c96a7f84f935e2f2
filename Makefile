# Builds libpointnerf_b200.so (sm_100a only) and the oracle's C port.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O3 -Xptxas -v
SRC := $(wildcard pointnerf_b200/csrc/*.cu)
OBJ := $(patsubst pointnerf_b200/csrc/%.cu,build/%.o,$(SRC))
LIB := pointnerf_b200/lib/libpointnerf_b200.so

all: $(LIB)

build/%.o: pointnerf_b200/csrc/%.cu $(wildcard pointnerf_b200/csrc/*.cuh) include/pointnerf_b200.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIB): $(OBJ)
	@mkdir -p pointnerf_b200/lib
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -lcuda

clean:
	rm -rf build $(LIB)
