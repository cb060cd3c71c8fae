# Top-level build: libdpsvm.so (CUDA, sm_100a), the host C++ layer and CLI, the oracle (checker).
#   make            everything
#   make lib        dpsvm_b200/libdpsvm.so only
NVCC      ?= nvcc
CXX       := g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unknown-pragmas
CSRC      := dpsvm_b200/csrc
OBJDIR    := build
LIB       := dpsvm_b200/libdpsvm.so
CU_SRCS   := $(CSRC)/smo_kernels.cu $(CSRC)/dpsvm_ctx.cu $(CSRC)/predict.cu $(CSRC)/comm.cu $(CSRC)/host_utils.cu $(CSRC)/noalloc.cu $(CSRC)/smo_resident.cu $(CSRC)/smo_sparse.cu
CU_OBJS   := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU_SRCS))
HDRS      := $(wildcard $(CSRC)/*.cuh) include/dpsvm_cuda.h

HOST      := dpsvm_b200/host
HOST_SRCS := $(HOST)/svmTrain.cpp $(HOST)/parse.cpp $(HOST)/model_io.cpp
HOST_LIB  := dpsvm_b200/libdpsvm_host.so
CSV_LIB   := dpsvm_b200/libdpsvm_csv.so
CLI       := dpsvm_b200/svmTrain

REF       ?= /root/reference

all: lib host oracle $(if $(wildcard $(REF)/svmTrainMain.cpp),dropin refgpu,)

# drop-in proof: the UNMODIFIED reference driver linked against this repo's SvmTrain (tools/dropin_check.sh)
dropin: host
	REF=$(REF) tools/dropin_check.sh

# the reference's own GPU path (cuBLAS gemv + Thrust) rebuilt for sm_100a from the sources where
# they lie, as a comparison binary for the same B200 (oracle/_ref is git-ignored and travels to the GPU box).
# -std=c++14: `beta` at svmTrain.cu:38 collides with std::beta under C++17; -include thrust/remove.h: svmTrain.cu:597.
refgpu: oracle/_ref/svmTrain_refgpu
oracle/_ref/svmTrain_refgpu: $(REF)/svmTrain.cu $(REF)/cache.cu $(REF)/parse.cpp $(REF)/svmTrainMain.cpp tools/dropin/mpi.h oracle/cblas_shim/cblas.h
	@mkdir -p oracle/_ref
	$(NVCC) $(ARCH) -x cu -O3 -std=c++14 -w -Ioracle/cblas_shim -Itools/dropin -include thrust/remove.h -o $@ \
	    $(REF)/svmTrain.cu $(REF)/cache.cu $(REF)/parse.cpp $(REF)/svmTrainMain.cpp -lcublas

lib: $(LIB)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(CU_OBJS) -lcudart -ldl

host: $(HOST_LIB) $(CLI) $(CSV_LIB)

# the CSV loader / writer alone (no SvmTrain, so no `extern state`): loadable with dlopen / ctypes
$(CSV_LIB): $(HOST)/parse.cpp include/parse.hpp
	$(CXX) -O2 -std=c++17 -fPIC -Wall -pthread -shared -Iinclude -o $@ $(HOST)/parse.cpp

$(HOST_LIB): $(HOST_SRCS) $(wildcard include/*.h include/*.hpp) $(LIB)
	$(CXX) -O2 -std=c++17 -fPIC -Wall -pthread -shared -Iinclude -o $@ $(HOST_SRCS) -Ldpsvm_b200 -ldpsvm -Wl,-rpath,'$$ORIGIN'

$(CLI): $(HOST)/svmTrainMain.cpp $(HOST_LIB)
	$(CXX) -O2 -std=c++17 -Wall -Iinclude -o $@ $(HOST)/svmTrainMain.cpp -Ldpsvm_b200 -ldpsvm_host -ldpsvm -Wl,-rpath,'$$ORIGIN'

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf $(OBJDIR) $(LIB) $(HOST_LIB) $(CSV_LIB) $(CLI)
	$(MAKE) -C oracle clean

.PHONY: all lib host oracle clean dropin refgpu
