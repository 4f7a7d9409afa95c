# Builds libelb200.so (device layer + host engine, sm_100a only) and the bring-up tools.
# `python -c "import __graft_entry__ as g; g.build()"` drives the same commands.
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := -std=c++17 -O3 -lineinfo $(ARCH) -Xcompiler -fPIC,-Wall,-Wno-unused-function
CSRC      := elemental_b200/csrc
LIBDIR    := elemental_b200/lib
CU_SRCS   := $(wildcard $(CSRC)/*.cu)
CPP_SRCS  := $(wildcard $(CSRC)/*.cpp)
OBJS      := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU_SRCS)) $(patsubst $(CSRC)/%.cpp,build/%.cpp.o,$(CPP_SRCS))
HDRS      := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh $(CSRC)/*.hpp include/*.h)
NCCL_INC  ?= /usr/include
NCCL_LIB  ?= $(shell python -c "import nvidia.nccl, os; print(os.path.join(list(nvidia.nccl.__path__)[0], 'lib'))" 2>/dev/null)

all: $(LIBDIR)/libelb200.so tools/gemm_check tools/el_gemm_driver tools/el_api_check tools/el_level3_check tools/el_comm_check tools/native_bench

build/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -I$(NCCL_INC) -c $< -o $@

build/%.cpp.o: $(CSRC)/%.cpp $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -I$(NCCL_INC) -x cu -c $< -o $@

$(LIBDIR)/libelb200.so: $(OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart -lnccl

tools/gemm_check: tools/gemm_check.cu $(LIBDIR)/libelb200.so include/elb200.h
	$(NVCC) $(NVFLAGS) -o $@ $< -L$(LIBDIR) -lelb200 -lcublas -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

# counterpart of the reference's tests/blas_like/Gemm.cpp over the C++ facade (include/El.hpp)
tools/el_gemm_driver: tools/el_gemm_driver.cpp include/El.hpp include/elb200.h $(LIBDIR)/libelb200.so
	$(NVCC) -std=c++17 -O2 -Xcompiler -Wall -Iinclude -o $@ $< -L$(LIBDIR) -lelb200 -lcudart -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

# every name of the facade, written the way a Hydrogen caller writes it (compile + link = the CPU-side check)
tools/el_api_check: tools/el_api_check.cpp include/El.hpp include/elb200.h $(LIBDIR)/libelb200.so
	$(NVCC) -std=c++17 -O1 -Xcompiler -Wall -Iinclude -o $@ $< -L$(LIBDIR) -lelb200 -lcudart -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

# the triangular rank-k family of the facade (Syrk / Herk / Trrk / LocalTrrk / Symm / Hemm, trapezoid routines)
tools/el_level3_check: tools/el_level3_check.cpp include/El.hpp include/elb200.h $(LIBDIR)/libelb200.so
	$(NVCC) -std=c++17 -O1 -Xcompiler -Wall -Iinclude -o $@ $< -L$(LIBDIR) -lelb200 -lcudart -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

# communicator side of the facade: Grid(mpi::COMM_WORLD, r), the grid's communicators, mpi:: collectives, proxies, MultiSync
tools/el_comm_check: tools/el_comm_check.cpp include/El.hpp include/elb200.h $(LIBDIR)/libelb200.so
	$(NVCC) -std=c++17 -O1 -Xcompiler -Wall -Iinclude -o $@ $< -L$(LIBDIR) -lelb200 -lcudart -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

# Python-free timings through the C ABI (e2e host-operand Gemm, fp64 kernel A/B vs cuBLAS, level-3 kernels)
tools/native_bench: tools/native_bench.cu $(LIBDIR)/libelb200.so include/elb200.h
	$(NVCC) $(NVFLAGS) -o $@ $< -L$(LIBDIR) -lelb200 -lcublas -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

clean:
	rm -rf build $(LIBDIR)/libelb200.so tools/gemm_check tools/el_gemm_driver tools/el_api_check tools/el_level3_check tools/el_comm_check tools/native_bench

.PHONY: all clean
