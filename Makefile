# Build of the native core for sm_100a (B200).  No GPU is needed to build.
#   make            -> game-of-life---mpi-cuda_b200/_lib/libgol_b200.so + bin/gol
#   make sass       -> profiles/sass/*.sass listings of the hot kernels
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function
PKG       := game-of-life---mpi-cuda_b200
LIB       := $(PKG)/_lib/libgol_b200.so
BIN       := bin/gol
OBJDIR    := build
CU_SRCS   := csrc/kernels.cu csrc/world.cu csrc/noalloc.cu
CXX_SRCS  := csrc/cpu_ref.cpp
CU_SRCS   := $(wildcard $(CU_SRCS))
OBJS      := $(patsubst csrc/%.cu,$(OBJDIR)/%.o,$(CU_SRCS)) $(patsubst csrc/%.cpp,$(OBJDIR)/%.o,$(CXX_SRCS))
HDRS      := $(wildcard csrc/*.h csrc/*.cuh include/*.h)

DBGLIB    := $(PKG)/_lib/libgol_b200_debug.so
DBGOBJS   := $(patsubst csrc/%.cu,$(OBJDIR)/debug/%.o,$(CU_SRCS)) $(patsubst csrc/%.cpp,$(OBJDIR)/debug/%.o,$(CXX_SRCS))

all: $(LIB) $(BIN) $(DBGLIB) bin/microbench_sync

# bounds-checking build: same sources with -DGOL_DEBUG_BOUNDS=1 (include/gol_b200.h, "Memory-safety build")
$(OBJDIR)/debug/%.o: csrc/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)/debug
	$(NVCC) $(NVFLAGS) -DGOL_DEBUG_BOUNDS=1 -c $< -o $@

$(OBJDIR)/debug/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(OBJDIR)/debug
	$(NVCC) $(NVFLAGS) -DGOL_DEBUG_BOUNDS=1 -x cu -c $< -o $@

$(DBGLIB): $(DBGOBJS)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $(DBGOBJS) -lcudart_static -ldl -lrt -lpthread

debug: $(DBGLIB)

$(OBJDIR)/%.o: csrc/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(OBJDIR)/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIB): $(OBJS)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart_static -ldl -lrt -lpthread

$(BIN): csrc/cli_main.cpp $(LIB) $(HDRS)
	@mkdir -p bin
	g++ -O2 -std=c++17 -Wall -o $@ csrc/cli_main.cpp -L$(PKG)/_lib -lgol_b200 -Wl,-rpath,'$$ORIGIN/../$(PKG)/_lib' -lpthread

# per-kernel SASS listings under profiles/sass/ (no GPU needed)
sass: $(LIB)
	python scripts/split_sass.py

# hand-off latencies between CTAs (cluster barrier, DSMEM, L2): run on a GPU box
bin/microbench_sync: scripts/microbench_sync.cu
	@mkdir -p bin
	$(NVCC) $(ARCH) -O3 -o $@ $<

microbench: bin/microbench_sync

test: all
	python -m pytest tests -q -m "not gpu"

test-gpu: all
	python -m pytest tests -q -m gpu

clean:
	rm -rf $(OBJDIR) $(LIB) $(DBGLIB) $(BIN)

.PHONY: all clean sass test test-gpu debug microbench
