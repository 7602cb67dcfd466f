# Top-level build of the product: libb200stream.so (sm_100a kernels + C ABI) and, when a BabelStream
# tree is available, b200-stream = the reference's UNMODIFIED src/main.cpp + our B200Stream<T> model.
#
#   make lib      babelstream_b200/csrc/libb200stream.so
#   make driver   build/b200-stream          (needs BABELSTREAM_ROOT, default /root/reference)
#   make oracle   the checker under oracle/  (never linked into the product)
#   make all      everything
#
# nvcc cross-compiles for sm_100a without a GPU.  The environment's CC/CXX wrappers cannot link
# OpenMP, hence the explicit /usr/bin host compiler.  Host code uses -march=x86-64-v3, not native:
# the artefacts are built in the CPU container and run on the GPU box.

BABELSTREAM_ROOT ?= /root/reference
NVCC    ?= /usr/local/cuda/bin/nvcc
HOSTCXX ?= /usr/bin/g++
MARCH   ?= x86-64-v3

CSRC    := babelstream_b200/csrc
MODEL   := babelstream_b200/src/b200
OBJDIR  := build/obj
LIB     := $(CSRC)/libb200stream.so

# -fmad=false: every FMA in the kernels is written by hand (rounding parity with the CPU reference)
NVCCFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false \
             -ccbin $(HOSTCXX) -Xcompiler -fPIC,-march=$(MARCH),-Wall -Iinclude

SRCS := $(wildcard $(CSRC)/*.cu)
OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(SRCS))
HDRS := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) include/b200_stream.h

.PHONY: all lib driver oracle clean ptxas probe lab sass
all: lib oracle driver

lib: $(LIB)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVCCFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; exit 1)
	@grep -E "error|warning" $(OBJDIR)/$*.ptxas.log | grep -v "^ptxas info" || true

$(LIB): $(OBJS)
	$(NVCC) -shared -ccbin $(HOSTCXX) -o $@ $^ -ldl -lpthread

oracle:
	$(MAKE) -C oracle all BABELSTREAM_ROOT=$(BABELSTREAM_ROOT)

# ---- b200-stream: reference driver + our model ------------------------------------------------
# main.cpp includes "StreamModels.h" with quotes, so the file next to it always wins over -I.
# The reference's main.cpp is therefore fed to the compiler on stdin, byte for byte, from where it
# lies: for stdin the "directory of the current file" is the working directory, build/driver/,
# which holds a generated StreamModels.h = the reference's text + the B200 branch
# (tools/gen_stream_models.py).  No copy of main.cpp is made.
ifneq ($(wildcard $(BABELSTREAM_ROOT)/src/main.cpp),)
ABS := $(abspath .)
driver: build/b200-stream
build/driver/StreamModels.h: $(BABELSTREAM_ROOT)/src/StreamModels.h tools/gen_stream_models.py
	@mkdir -p build/driver
	python tools/gen_stream_models.py $< $@
build/driver/main.o: build/driver/StreamModels.h $(BABELSTREAM_ROOT)/src/main.cpp $(MODEL)/B200Stream.h include/b200_stream.h
	cd build/driver && $(HOSTCXX) -std=c++14 -O3 -march=$(MARCH) -DB200 -I. -I$(BABELSTREAM_ROOT)/src \
	    -I$(ABS)/$(MODEL) -I$(ABS)/include -x c++ -c -o main.o - < $(BABELSTREAM_ROOT)/src/main.cpp
	sha256sum $(BABELSTREAM_ROOT)/src/main.cpp > build/driver/main.cpp.sha256
build/b200-stream: build/driver/main.o $(MODEL)/B200Stream.cpp $(MODEL)/B200Stream.h $(LIB)
	$(HOSTCXX) -std=c++14 -O3 -march=$(MARCH) -DB200 -I$(BABELSTREAM_ROOT)/src -I$(MODEL) -Iinclude \
	    build/driver/main.o $(MODEL)/B200Stream.cpp -o $@ \
	    -L$(CSRC) -lb200stream -Wl,-rpath,'$$ORIGIN/../$(CSRC)' -lpthread -ldl
else
driver:
	@echo "Makefile: $(BABELSTREAM_ROOT)/src/main.cpp not present -- b200-stream not rebuilt"
endif

# tools/latency_probe: host-clock split of a tiny blocking call (bare kernel vs LEVEL 1 vs LEVEL 2)
probe: build/latency_probe
build/latency_probe: tools/latency_probe.cu $(LIB) include/b200_stream.h
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -ccbin $(HOSTCXX) -Iinclude $< -o $@ \
	    -L$(CSRC) -lb200stream -Xlinker -rpath,'$$ORIGIN/../$(CSRC)'

ptxas:
	@cat $(OBJDIR)/*.ptxas.log | grep -E "Compiling|registers|spill" | paste - - - | sort | uniq | head -2000

# opcode histograms of the default-selected kernels (and the async-proxy copy variants): the committed SASS evidence
sass: lib
	@mkdir -p profiles/r03
	python tools/sass_summary.py > profiles/r03/sass_summary.txt
	@echo wrote profiles/r03/sass_summary.txt

clean:
	rm -rf build $(LIB)

# tools/lab/ew_lab: experimental bench of elementwise-kernel candidates (not part of the product)
lab: build/ew_lab
build/ew_lab: tools/lab/ew_lab.cu
	@mkdir -p build
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -ccbin $(HOSTCXX) --threads 0 -o $@ $<
