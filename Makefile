# gol-b200 build.  Everything is built in-tree for sm_100a (no install step).
#
#   make            library (libgolb200.so) + bin/gol + bin/gol-generate + the micro-benchmarks
#                   (bin/gol-intpeak: integer-pipe peak; bin/gol-tilebench: TMA / cluster tile-kernel experiment)
#   make game | mpi | collective | async | openmp | cuda
#                   the reference's six make targets (Makefile:12-28 of the reference): each
#                   writes ./a.out with that variant's report format and output file name.
#   make oracle     builds the unmodified reference serial program into baseline/_ref (tests)
#   make clean      removes built objects only (the reference's `clean` also deletes the
#                   output grids because it removes *.out, defect D11; this one does not).
NVCC      ?= nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v $(EXTRA_NVFLAGS)
CXXFLAGS  := -O3 -std=c++17 -fPIC -Wall -pthread
BUILD     := build
LIB       := libgolb200.so
CU_SRCS   := csrc/life_kernels.cu csrc/dense_kernels.cu csrc/engine.cu csrc/static_scratch.cu
CPP_SRCS  := csrc/host_util.cpp
OBJS      := $(CU_SRCS:csrc/%.cu=$(BUILD)/%.o) $(CPP_SRCS:csrc/%.cpp=$(BUILD)/%.o)
HDRS      := include/gol_b200.h csrc/gol_internal.h csrc/life_rule.h csrc/life_march.cuh csrc/life_items.h

all: $(LIB) bin/gol bin/gol-generate bin/gol-intpeak bin/gol-tilebench

$(BUILD)/%.o: csrc/%.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(BUILD)/$*.ptxas.log || (cat $(BUILD)/$*.ptxas.log; false)

$(BUILD)/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -Xlinker -soname=$(LIB) -ldl

lib: $(LIB)

bin/gol: csrc/cli_main.cpp $(LIB)
	@mkdir -p bin
	$(CXX) $(CXXFLAGS) -DGOL_DEFAULT_VARIANT='"game"' $< -o $@ -L. -lgolb200 -Wl,-rpath,'$$ORIGIN/..' -Wl,-rpath,'$$ORIGIN'

bin/gol-generate: csrc/generate_main.cpp $(LIB)
	@mkdir -p bin
	$(CXX) $(CXXFLAGS) $< -o $@ -L. -lgolb200 -Wl,-rpath,'$$ORIGIN/..' -Wl,-rpath,'$$ORIGIN'

bin/gol-intpeak: csrc/int_peak.cu
	@mkdir -p bin
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo $< -o $@

# experiment (not linked into the library): TMA-fed cluster/DSMEM tile kernel against the product's march core
bin/gol-tilebench: csrc/experiments/life_tile.cu csrc/life_march.cuh csrc/gol_internal.h include/gol_b200.h
	@mkdir -p bin $(BUILD)
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo -Xptxas -v -Iinclude $< -o $@ 2> $(BUILD)/life_tile.ptxas.log || (cat $(BUILD)/life_tile.ptxas.log; false)

# the reference's make targets: same binary, different default variant, written to ./a.out
game mpi collective async openmp cuda: $(LIB)
	$(CXX) $(CXXFLAGS) -DGOL_DEFAULT_VARIANT='"$@"' csrc/cli_main.cpp -o a.out -L. -lgolb200 -Wl,-rpath,'$$ORIGIN'

oracle:
	$(MAKE) -C baseline ref_serial

clean:
	rm -rf $(BUILD) $(LIB) bin/gol bin/gol-generate bin/gol-intpeak bin/gol-tilebench a.out

.PHONY: all lib game mpi collective async openmp cuda oracle clean
