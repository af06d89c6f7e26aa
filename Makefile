# Top-level build: the product (librtb200.so = hand-written sm_100a CUDA behind a C ABI, the
# C++ host scene API on top of it, the headless CLI) and the checkers under oracle/.
# Everything is built IN-TREE so that the binaries travel with the snapshot to the GPU box.
ROOT    := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))
PKG     := $(ROOT)ray-tracing-gpu_b200
CSRC    := $(PKG)/csrc
HOST    := $(PKG)/host
LIBDIR  := $(PKG)/lib
OBJDIR  := $(PKG)/build
NVCC    ?= /usr/local/cuda/bin/nvcc
HOSTCXX := /usr/bin/g++

NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -ccbin $(HOSTCXX) \
           -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -warn-spills
CUOBJS  := $(addprefix $(OBJDIR)/,rt_api.o rt_resident.o rt_stream.o rt_trace.o rt_comm.o rt_peak.o rt_bvh.o rt_sequence.o rt_pack.o rt_cullscan.o rt_noalloc.o)
CUHDRS  := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) $(ROOT)include/rtb200.h

all: lib host oracle

lib: $(LIBDIR)/librtb200.so

$(OBJDIR)/%.o: $(CSRC)/%.cu $(CUHDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIBDIR)/librtb200.so: $(CUOBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) -shared -ccbin $(HOSTCXX) -o $@ $^ -ldl

# C++ host scene API (fresh implementation of the reference's CScene / ISurface hierarchy)
# and the headless CLI that replaces Main.cpp's GLUT window.
HOSTSRC := $(wildcard $(HOST)/src/*.cpp)
HOSTHDR := $(wildcard $(HOST)/include/*.h) $(wildcard $(HOST)/src/*.hpp)
HOSTFLAGS := -std=c++17 -O2 -ffp-contract=off -fPIC -Wall -pthread -I$(HOST)/include -I$(HOST)/src -I$(ROOT)include

host: $(LIBDIR)/librtb200_scene.so $(LIBDIR)/rt_cli

$(LIBDIR)/librtb200_scene.so: $(filter-out $(HOST)/src/rt_cli.cpp,$(HOSTSRC)) $(HOSTHDR) $(LIBDIR)/librtb200.so
	$(HOSTCXX) $(HOSTFLAGS) -shared -o $@ $(filter-out $(HOST)/src/rt_cli.cpp,$(HOSTSRC)) -L$(LIBDIR) -lrtb200 -Wl,-rpath,'$$ORIGIN'

$(LIBDIR)/rt_cli: $(HOST)/src/rt_cli.cpp $(LIBDIR)/librtb200_scene.so
	$(HOSTCXX) $(HOSTFLAGS) -o $@ $(HOST)/src/rt_cli.cpp -L$(LIBDIR) -lrtb200_scene -lrtb200 -Wl,-rpath,'$$ORIGIN'

oracle: lib
	$(MAKE) -C $(ROOT)oracle all

# CHECKED build: every .cu with -DRTB200_CHECKED (device-side bounds checks, rt_math.cuh: RT_CHECK) into ab/checked/
# (git-ignored, travels to the GPU box); run anything against it with RTB200_LIB=ab/checked/librtb200.so
CHKDIR  := $(ROOT)ab/checked
CHKOBJS := $(patsubst $(OBJDIR)/%.o,$(CHKDIR)/%.o,$(CUOBJS))
$(CHKDIR)/%.o: $(CSRC)/%.cu $(CUHDRS)
	@mkdir -p $(CHKDIR)
	$(NVCC) $(NVFLAGS) -DRTB200_CHECKED -c $< -o $@
checked: $(CHKOBJS)
	$(NVCC) -shared -ccbin $(HOSTCXX) -o $(CHKDIR)/librtb200.so $^ -ldl

# ASan + UBSan build of the host scene API (parser, Pretraitement, SoA packer) and of the C oracle into ab/asan/; run the
# CPU tests against them with
#   RTB200_SCENE_LIB=ab/asan/librtb200_scene.so RTB200_ORACLE_LIB=ab/asan/librt_oracle.so \
#   LD_PRELOAD="$$(gcc -print-file-name=libasan.so) $$(gcc -print-file-name=libubsan.so)" \
#   ASAN_OPTIONS=detect_leaks=0:alloc_dealloc_mismatch=0:new_delete_type_mismatch=0 python -m pytest tests -m "not gpu"
# (the two mismatch checks are off because the REFERENCE frees its new[] frame buffer with a scalar delete, Scene.cpp:528)
asan: $(LIBDIR)/librtb200.so
	@mkdir -p $(ROOT)ab/asan
	$(HOSTCXX) $(HOSTFLAGS) -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -shared -o $(ROOT)ab/asan/librtb200_scene.so \
	    $(filter-out $(HOST)/src/rt_cli.cpp,$(HOSTSRC)) -L$(LIBDIR) -lrtb200 -Wl,-rpath,$(LIBDIR)
	/usr/bin/gcc -std=gnu11 -O1 -g -fsanitize=address,undefined -ffp-contract=off -fPIC -fopenmp -shared -o $(ROOT)ab/asan/librt_oracle.so $(ROOT)oracle/rt_oracle.c -lm

clean:
	rm -rf $(OBJDIR) $(LIBDIR)
	$(MAKE) -C $(ROOT)oracle clean

.PHONY: all lib host oracle checked asan clean
