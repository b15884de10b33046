# Builds the product library mundy_b200/libmundy_b200.so (sm_100a only) and, via oracle/Makefile, the test oracle.
# -fmad=false: fp64 results follow the operation order written in the kernels (bit-level parity with the reference's
# CPU evaluation); every stage is HBM-bound so contraction would buy nothing.
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
             -Xcompiler -Wall -Xptxas -v
SRCDIR    := mundy_b200/csrc
SRCS      := $(wildcard $(SRCDIR)/*.cu)
OBJS      := $(patsubst $(SRCDIR)/%.cu,build/%.o,$(SRCS))
HDRS      := $(wildcard $(SRCDIR)/*.cuh) include/mundy_b200.h
LIB       := mundy_b200/libmundy_b200.so

all: $(LIB)

build/%.o: $(SRCDIR)/%.cu $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build $(LIB)
.PHONY: all oracle clean
