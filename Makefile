# Same user-facing targets as the reference build (reference: Makefile:4-15):
#   make all | engine | engine.debug | clean
# plus the shared libraries the Python package loads and a few tools.
#   make NO_CUDA=1 ...   builds the CPU-only driver (no nvcc needed)
CXX      ?= g++
NVCC     ?= /usr/local/cuda/bin/nvcc
CUDA_HOME ?= /usr/local/cuda
CXXFLAGS ?= -O3 -std=c++17 -ffp-contract=off -fPIC -Wall -Wextra -pthread -Iinclude
SM_FLAGS  = -gencode arch=compute_100a,code=sm_100a
NVFLAGS  ?= -O3 -std=c++17 $(SM_FLAGS) -lineinfo -Xcompiler -fPIC,-ffp-contract=off,-pthread -Iinclude --fmad=true

HOST_SRCS = src/driver/main.cpp src/driver/report.cpp src/driver/io.cpp \
            src/engine/engine.cpp src/cpu/cpu_backends.cpp
KERNEL_SRCS = $(wildcard src/kernels/*.cu)
RUNTIME_SRCS = $(wildcard src/runtime/*.cu) $(wildcard src/engine/*.cu)
LIBDIR = distributed-machine-learning-project_b200/lib

.PHONY: all clean libs tools test

all: engine engine.debug

ifeq ($(NO_CUDA),1)
engine: $(HOST_SRCS) src/engine/gpu_stub.cpp
	$(CXX) $(CXXFLAGS) $^ -o $@
engine.debug: $(HOST_SRCS) src/engine/gpu_stub.cpp
	$(CXX) $(CXXFLAGS) -g -DDEBUG $^ -o $@
else
engine: $(HOST_SRCS) $(KERNEL_SRCS) $(RUNTIME_SRCS)
	$(NVCC) $(NVFLAGS) $^ -o $@ -lcudart
engine.debug: $(HOST_SRCS) $(KERNEL_SRCS) $(RUNTIME_SRCS)
	$(NVCC) $(NVFLAGS) -g -DDEBUG $^ -o $@ -lcudart
endif

# Shared objects loaded by the Python package through ctypes.
libs: $(LIBDIR)/libdmlp_cpu.so $(LIBDIR)/libdmlp_knn.so

$(LIBDIR)/libdmlp_cpu.so: src/cpu/cpu_backends.cpp src/driver/io.cpp src/cpu/c_api.cpp
	mkdir -p $(LIBDIR)
	$(CXX) $(CXXFLAGS) -shared $^ -o $@

$(LIBDIR)/libdmlp_knn.so: $(KERNEL_SRCS) $(wildcard src/runtime/*.cu)
	mkdir -p $(LIBDIR)
	$(NVCC) $(NVFLAGS) -shared $^ -o $@ -lcudart

tools: tools/dmlp-gen
tools/dmlp-gen: tools/dmlp_gen.cpp src/driver/io.cpp
	$(CXX) $(CXXFLAGS) $^ -o $@

clean:
	rm -rf engine engine.debug $(LIBDIR)/*.so tools/dmlp-gen outputs/tmp.*
