# Top-level build: product libraries (CUDA, sm_100a only), host application, oracle.
#   make            -> everything
#   make product    -> owl_path_tracer_b200/lib/libb200pt.so, libb200pt_host.so, bin/path_tracer
#   make oracle     -> oracle/liboracle.so, oracle/liboracle_det.so   (test infrastructure)
#   make emu        -> tests/emu/libemu.so  (host emulation of the kernels' per-lane logic; tests only)
#   make ref        -> oracle/_ref/libref_device.so, libref_host.so: the REFERENCE'S OWN device.cu / disney/*.cuh /
#                      loaders compiled for the host from /root/reference (tests only; no-op where it is absent)
NVCC ?= /usr/local/cuda/bin/nvcc
# the environment may export CXX=/opt/gcc/bin/g++, a wrapper that links libstdc++ statically: several such copies in one
# Python process (next to torch's dynamic one) break exception handling, so the system compiler is named explicitly
CXX = /usr/bin/g++
PKG = owl_path_tracer_b200
CSRC = $(PKG)/csrc
HOST = $(PKG)/host
LIB = $(PKG)/lib
BIN = $(PKG)/bin

NVFLAGS = -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false \
          -Xcompiler -fPIC -Xcompiler -ffp-contract=off --expt-relaxed-constexpr -cudart static
CXXFLAGS = -std=c++17 -O2 -fPIC -Wall -Wno-unused-function -pthread

CUHDR = $(wildcard $(CSRC)/*.h) $(wildcard $(CSRC)/*.cuh) include/b200pt.h
HOSTHDR = $(wildcard $(HOST)/*.hpp) include/b200pt.h include/b200pt_host.h

all: product oracle emu ref
product: $(LIB)/libb200pt.so $(LIB)/libb200pt_scene.so $(LIB)/libb200pt_host.so $(BIN)/path_tracer

$(LIB)/%.o: $(CSRC)/%.cu $(CUHDR)
	@mkdir -p $(LIB)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c -o $@ $< 2> $(LIB)/$*.ptxas.log || (cat $(LIB)/$*.ptxas.log; exit 1)

$(LIB)/libb200pt.so: $(LIB)/api.o $(LIB)/wavefront.o $(LIB)/bvh_build.o $(LIB)/mesh_dedup.o
	$(NVCC) -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o $@ $^

# libb200pt_scene.so: loaders + scene stages, NO CUDA dependency (CPU-only processes load scenes through it);
# libb200pt_host.so: bind / render / test_loop on top of it and of libb200pt.so
SCENESRC = $(HOST)/loaders.cpp $(HOST)/procedural.cpp $(HOST)/scene_load.cpp $(HOST)/scene_api.cpp
$(LIB)/libb200pt_scene.so: $(SCENESRC) $(HOSTHDR)
	@mkdir -p $(LIB)
	$(CXX) $(CXXFLAGS) -shared -o $@ $(SCENESRC) -lz

HOSTSRC = $(HOST)/application.cpp $(HOST)/host_api.cpp
$(LIB)/libb200pt_host.so: $(HOSTSRC) $(HOSTHDR) $(LIB)/libb200pt.so $(LIB)/libb200pt_scene.so
	$(CXX) $(CXXFLAGS) -shared -o $@ $(HOSTSRC) -L$(LIB) -lb200pt_scene -lb200pt -Wl,-rpath,'$$ORIGIN'

$(BIN)/path_tracer: $(HOST)/main.cpp $(LIB)/libb200pt_host.so
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) -fPIE -o $@ $(HOST)/main.cpp -L$(LIB) -lb200pt_host -lb200pt_scene -lb200pt -Wl,-rpath,'$$ORIGIN/../lib'

oracle:
	$(MAKE) -C oracle

ref:
	$(MAKE) -C oracle/ref

emu: tests/emu/libemu.so
tests/emu/libemu.so: tests/emu/emu.cpp $(CUHDR)
	$(CXX) -std=c++17 -O2 -fPIC -shared -ffp-contract=off -mfma -fno-fast-math -Wall -Wno-unused-function -Wno-unknown-pragmas -pthread -o $@ $<

clean:
	rm -rf $(LIB) $(BIN) tests/emu/libemu.so
	$(MAKE) -C oracle clean
	$(MAKE) -C oracle/ref clean

.PHONY: all product oracle emu ref clean
