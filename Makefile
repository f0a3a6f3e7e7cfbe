# Same user-facing targets as the reference build (reference: Makefile:4-15):
#   make all | engine | engine.debug | clean
# plus the shared libraries the Python package loads (make libs) and tools.
#   make NO_CUDA=1 ...   builds a CPU-only driver (no nvcc needed)
CXX       ?= g++
NVCC      ?= /usr/local/cuda/bin/nvcc
CXXFLAGS  ?= -O3 -std=c++17 -ffp-contract=off -fPIC -Wall -Wextra -pthread -Iinclude -MMD -MP
SM_FLAGS   = -gencode arch=compute_100a,code=sm_100a
NVFLAGS   ?= -O3 -std=c++17 --expt-relaxed-constexpr $(SM_FLAGS) -lineinfo \
             -Xcompiler -fPIC,-ffp-contract=off,-pthread -Iinclude -MMD -MP

B       = build
LIBDIR  = distributed-machine-learning-project_b200/lib

HOST_CPP   = src/driver/report.cpp src/driver/io.cpp src/engine/engine.cpp src/cpu/cpu_backends.cpp
KERNEL_CU  = $(wildcard src/kernels/*.cu) $(wildcard src/runtime/*.cu)
ENGINE_CU  = src/engine/gpu_engine.cu src/engine/multicast.cu

HOST_O     = $(HOST_CPP:%.cpp=$(B)/%.o)
HOST_DBG_O = $(HOST_CPP:%.cpp=$(B)/%.dbg.o)
KERNEL_O   = $(KERNEL_CU:%.cu=$(B)/%.o)
ENGINE_O   = $(ENGINE_CU:%.cu=$(B)/%.o)

.PHONY: all clean libs tools
all: engine engine.debug

$(B)/%.o: %.cpp
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -c $< -o $@
$(B)/%.dbg.o: %.cpp
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -g -DDEBUG -c $< -o $@
$(B)/%.o: %.cu
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c $< -o $@

ifeq ($(NO_CUDA),1)
engine: $(HOST_O) $(B)/src/driver/main.o $(B)/src/engine/gpu_stub.o
	$(CXX) -pthread $^ -o $@
engine.debug: $(HOST_DBG_O) $(B)/src/driver/main.dbg.o $(B)/src/engine/gpu_stub.dbg.o
	$(CXX) -pthread $^ -o $@
else
engine: $(HOST_O) $(B)/src/driver/main.o $(KERNEL_O) $(ENGINE_O)
	$(NVCC) $(SM_FLAGS) $^ -o $@ -lcudart -lpthread
engine.debug: $(HOST_DBG_O) $(B)/src/driver/main.dbg.o $(KERNEL_O) $(ENGINE_O)
	$(NVCC) $(SM_FLAGS) $^ -o $@ -lcudart -lpthread
endif

# Shared objects loaded by the Python package through ctypes.
libs: $(LIBDIR)/libdmlp_cpu.so $(LIBDIR)/libdmlp_knn.so

$(LIBDIR)/libdmlp_cpu.so: $(B)/src/cpu/cpu_backends.o $(B)/src/driver/io.o $(B)/src/cpu/c_api.o
	@mkdir -p $(LIBDIR)
	$(CXX) -shared -pthread $^ -o $@

$(LIBDIR)/libdmlp_knn.so: $(KERNEL_O)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(SM_FLAGS) -shared $^ -o $@ -lcudart

# ATen operator library (torch.ops.dmlp_b200.*): the same kernels behind PyTorch's dispatcher,
# current stream and caching allocator.  Built against the torch of $(PYTHON).
PYTHON    ?= python
TORCH_DIR  = $(shell $(PYTHON) -c "import os, torch; print(os.path.dirname(torch.__file__))" 2>/dev/null)
TORCH_ABI  = $(shell $(PYTHON) -c "import torch; print(int(torch._C._GLIBCXX_USE_CXX11_ABI))" 2>/dev/null)
.PHONY: torchops
torchops: $(LIBDIR)/libdmlp_torch.so
$(LIBDIR)/libdmlp_torch.so: src/torch/ops.cpp include/dmlp_knn.h $(LIBDIR)/libdmlp_knn.so
	$(CXX) -O2 -std=c++17 -fPIC -shared -Wall -D_GLIBCXX_USE_CXX11_ABI=$(TORCH_ABI) -Iinclude \
	  -isystem $(TORCH_DIR)/include -isystem $(TORCH_DIR)/include/torch/csrc/api/include -isystem /usr/local/cuda/include \
	  $< -o $@ -L$(LIBDIR) -ldmlp_knn -L$(TORCH_DIR)/lib -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda \
	  -Wl,-rpath,'$$ORIGIN' -Wl,-rpath,$(TORCH_DIR)/lib

tools: tools/dmlp-gen
tools/dmlp-gen: $(B)/tools/dmlp_gen.o $(B)/src/driver/io.o
	$(CXX) -pthread $^ -o $@

clean:
	rm -rf $(B) engine engine.debug $(LIBDIR)/*.so tools/dmlp-gen outputs/tmp.*

-include $(shell find $(B) -name '*.d' 2>/dev/null)

# ---- sanitizers (the reference has none) --------------------------------------
# Host: CPU-only driver under ASan + UBSan, run over every CPU strategy on a generated input.
.PHONY: asan sanitize-gpu
ASAN_CXX ?= /usr/bin/g++
ASAN_FLAGS = -O1 -g -std=c++17 -ffp-contract=off -pthread -Iinclude -fsanitize=address,undefined -fno-omit-frame-pointer
engine.asan: $(HOST_CPP) src/driver/main.cpp src/engine/gpu_stub.cpp
	$(ASAN_CXX) $(ASAN_FLAGS) $^ -o $@
asan: engine.asan
	python3 tools/generate_input.py --num_data 600 --num_queries 50 --num_attrs 7 --min -5 --max 5 --minK 1 --maxK 9 \
	    --num_labels 4 --output /tmp/dmlp_asan.in > /dev/null
	for s in serial grid gather tree farm kdtree; do \
	  DMLP_BACKEND=cpu DMLP_STRATEGY=$$s DMLP_RANKS=6 ./engine.asan < /tmp/dmlp_asan.in | sort | md5sum || exit 1; done
	DMLP_BACKEND=cpu DMLP_STRATEGY=grid DMLP_GRID=4x2 ./engine.asan --aos < /tmp/dmlp_asan.in | sort | md5sum
# Device: every kernel on small shapes under compute-sanitizer (one tool per run: TOOL=memcheck|racecheck|synccheck).
TOOL ?= memcheck
sanitize-gpu: libs
	compute-sanitizer --tool $(TOOL) --error-exitcode 9 python3 scripts/sanitize_case.py

# ---- examples: plain C against the C ABI (needs `make libs`) -----------------------------------
CUDA_HOME ?= /usr/local/cuda
examples: examples/knn_local
examples/knn_local: examples/knn_local.c include/dmlp_knn.h $(LIBDIR)/libdmlp_knn.so
	$(CC) -std=c99 -O2 -Wall -Wextra -Iinclude -I$(CUDA_HOME)/include $< -o $@ \
	    -L$(LIBDIR) -ldmlp_knn -L$(CUDA_HOME)/lib64 -lcudart -Wl,-rpath,'$$ORIGIN/../$(LIBDIR)' -Wl,-rpath,$(CUDA_HOME)/lib64
.PHONY: examples
