# Builds the product library (CUDA kernels + C++ host layer + C API) in-tree for sm_100a.
#   make            -> dla-future_b200/lib/libdlaf_b200.so
#   make miniapp    -> dla-future_b200/bin/miniapp_{cholesky,triangular_solver,gen_to_std}
#   make oracle     -> oracle/_build/liboracle.so (test infrastructure only)
#   make emu        -> tests/emu/_build/libsolver_emu.so (test infrastructure only)
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
PKG       := dla-future_b200
CSRC      := $(PKG)/csrc
SRC       := $(PKG)/src
BUILD     := build
LIBDIR    := $(PKG)/lib
BINDIR    := $(PKG)/bin
NVFLAGS   := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function -Iinclude -I$(CSRC) -I$(SRC)

CU_SRCS   := $(wildcard $(CSRC)/*.cu) $(wildcard $(SRC)/*.cu)
OBJS      := $(patsubst %.cu,$(BUILD)/%.o,$(notdir $(CU_SRCS)))
HDRS      := $(wildcard $(CSRC)/*.cuh) $(wildcard $(SRC)/*.h) $(wildcard include/*.h) $(wildcard include/*/*.h) \
             $(wildcard include/*/*/*.h) $(wildcard include/*/*/*/*.h)

vpath %.cu $(CSRC) $(SRC)

all: $(LIBDIR)/libdlaf_b200.so

$(BUILD)/%.o: %.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIBDIR)/libdlaf_b200.so: $(OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart -lnccl

miniapp: $(BINDIR)/miniapp_cholesky $(BINDIR)/miniapp_triangular_solver $(BINDIR)/miniapp_gen_to_std

$(BINDIR)/miniapp_triangular_solver $(BINDIR)/miniapp_gen_to_std: $(BINDIR)/%: miniapp/%.cpp miniapp/miniapp_common.h $(LIBDIR)/libdlaf_b200.so $(HDRS)
	@mkdir -p $(BINDIR)
	$(NVCC) $(ARCH) -O2 -std=c++17 -Iinclude -Iinclude/dlafb_mpi_shim -o $@ $< -L$(LIBDIR) -ldlaf_b200 -Xlinker -rpath -Xlinker '$$ORIGIN/../lib' -lcudart

$(BINDIR)/miniapp_cholesky: miniapp/miniapp_cholesky.cpp $(LIBDIR)/libdlaf_b200.so $(HDRS)
	@mkdir -p $(BINDIR)
	$(NVCC) $(ARCH) -O2 -std=c++17 -Iinclude -Iinclude/dlafb_mpi_shim -o $@ $< -L$(LIBDIR) -ldlaf_b200 -Xlinker -rpath -Xlinker '$$ORIGIN/../lib' -lcudart

oracle:
	@mkdir -p oracle/_build
	g++ -O2 -fopenmp -shared -fPIC -o oracle/_build/liboracle.so oracle/chol_oracle.cpp -ldl

# CPU execution of the triangular solver's algorithm template and kernel bodies (test infrastructure only)
emu: tests/emu/_build/libsolver_emu.so tests/emu/_build/libsimt_wide_emu.so

tests/emu/_build/libsimt_wide_emu.so: tests/emu/simt_wide_emu.cu $(CSRC)/gemm_simt_wide.cuh $(CSRC)/gemm_simt.cuh $(CSRC)/common.cuh
	@mkdir -p tests/emu/_build
	$(NVCC) -O1 -std=c++17 -Wno-deprecated-gpu-targets -diag-suppress 20014 -shared -Xcompiler -fPIC,-pthread,-ffp-contract=off -I$(CSRC) -o $@ $<

tests/emu/_build/libsolver_emu.so: tests/emu/solver_emu.cu $(SRC)/triangular_solver.h $(SRC)/gen_to_std.h $(SRC)/triangular_solver_dist.h $(SRC)/gen_to_std_dist.h $(SRC)/redistribute.h $(CSRC)/copy_rects.cuh $(SRC)/schedule.h $(CSRC)/solver_kernels.cuh $(CSRC)/common.cuh
	@mkdir -p tests/emu/_build
	$(NVCC) -O2 -std=c++17 -Wno-deprecated-gpu-targets -shared -Xcompiler -fPIC,-pthread -I$(CSRC) -I$(SRC) -o $@ $<

clean:
	rm -rf $(BUILD) $(LIBDIR) $(BINDIR) tests/emu/_build

.PHONY: all miniapp oracle emu clean
