# Builds libbbb_b200.so (hand-written CUDA for sm_100a behind the C ABI of include/bbb_b200.h),
# the CPU oracle (test infrastructure) and the command-line tools.
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v
CSRC      := basebandboard_b200/csrc
LIB       := basebandboard_b200/libbbb_b200.so
SRCS      := $(wildcard $(CSRC)/*.cu)
OBJS      := $(patsubst $(CSRC)/%.cu,build/%.o,$(SRCS))
HDRS      := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh $(CSRC)/*.inc) include/bbb_b200.h

all: $(LIB) oracle cli/rnghunt

build/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -Xlinker -soname,libbbb_b200.so

oracle:
	$(MAKE) -C oracle

# drop-in for the reference's rnghunt binary (software/rnghunt/src/bin/rnghunt.rs), C++ over the C ABI
cli/rnghunt: cli/rnghunt.cpp include/bbb_b200.h $(LIB)
	g++ -O2 -std=c++17 -Wall -pthread -o $@ cli/rnghunt.cpp -Lbasebandboard_b200 -lbbb_b200 -Wl,-rpath,'$$ORIGIN/../basebandboard_b200'

clean:
	rm -rf build $(LIB) cli/rnghunt
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
