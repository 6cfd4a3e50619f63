# Build of the B200-native library, the host driver and the test-side helpers.
#   make            libnsgpu.so (sm_100a kernels + C ABI + host Grid/FluidSolver) and bin/main
#   make oracle     CPU oracle (test infrastructure): oracle/_ref/* from the reference's sources, when /root/reference exists
#   make emul       tests/emul/libnsemul.so: host compile of the per-cell device functions, for the CPU-only tests
NVCC ?= nvcc
CXX = /usr/bin/g++
PKG = navierstokessolver_b200
CSRC = $(PKG)/csrc
HOST = $(PKG)/host
NVFLAGS = -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC $(if $(NSGPU_BUILD_STAMPS),-DNSGPU_STAMPS=1)
CU_DEPS = $(wildcard $(CSRC)/*.cu $(CSRC)/*.cuh $(CSRC)/*.inl $(CSRC)/*.h) include/nsgpu.h
HOST_SRC = $(wildcard $(HOST)/*.cpp)
HOST_DEPS = $(wildcard $(HOST)/*.cpp $(HOST)/*.h)

.PHONY: all oracle emul clean
all: $(PKG)/libnsgpu.so bin/main

$(PKG)/libnsgpu.so: $(CU_DEPS) $(HOST_DEPS)
	$(NVCC) $(NVFLAGS) -shared $(CSRC)/nsgpu.cu $(HOST_SRC) -Iinclude -I$(HOST) -o $@ -lnccl -ldl

bin/main: $(PKG)/libnsgpu.so $(HOST)/main/ns_main.cpp
	@mkdir -p bin
	$(CXX) -std=c++17 -O2 -Iinclude -I$(HOST) $(HOST)/main/ns_main.cpp -L$(PKG) -lnsgpu -Wl,-rpath,'$$ORIGIN/../$(PKG)' -o $@

oracle:
	if [ -d /root/reference/SRC ]; then $(MAKE) -C oracle ref; fi

emul: tests/emul/libnsemul.so
tests/emul/libnsemul.so: tests/emul/emul.cpp $(CSRC)/ns_cell.cuh $(CSRC)/ns_hostgeom.h include/nsgpu.h
	$(CXX) -std=c++17 -O2 -fPIC -shared -Iinclude -I$(CSRC) tests/emul/emul.cpp -o $@

clean:
	rm -f $(PKG)/libnsgpu.so bin/main tests/emul/libnsemul.so
	$(MAKE) -C oracle clean
