# Builds libmpcb.so (the C-ABI product library: CUDA kernels for sm_100a + host planner code), in-tree.
NVCC ?= nvcc
CXX ?= g++
CUDA_HOME ?= /usr/local/cuda
SRC = mpc_motion_planner_b200/csrc
OUT = mpc_motion_planner_b200/libmpcb.so
NVFLAGS = -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Iinclude -I$(SRC) -Xcompiler -fPIC
CXXFLAGS = -O2 -std=c++17 -fPIC -Wall -Iinclude -I$(SRC) -I$(CUDA_HOME)/include
HDR = $(wildcard $(SRC)/*.h) include/mpcb.h

all: $(OUT)
build/mpcb_kernels.o: $(SRC)/mpcb_kernels.cu $(HDR)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@
build/mpcb_capi.o: $(SRC)/mpcb_capi.cpp $(HDR)
	@mkdir -p build
	$(CXX) $(CXXFLAGS) -c $< -o $@
build/mpcb_planner_host.o: $(SRC)/mpcb_planner_host.cpp $(HDR)
	@mkdir -p build
	$(CXX) $(CXXFLAGS) -c $< -o $@
$(OUT): build/mpcb_kernels.o build/mpcb_capi.o build/mpcb_planner_host.o
	$(NVCC) -shared -o $@ $^ -cudart shared -Xlinker -rpath=$(CUDA_HOME)/lib64
clean:
	rm -rf build $(OUT)

# Debug variant with per-phase cycle counters in the ADMM kernel (tools/phase_profile.py); not used by tests or bench.
PROF_OUT = mpc_motion_planner_b200/libmpcb_prof.so
prof: $(PROF_OUT)
build/prof_kernels.o: $(SRC)/mpcb_kernels.cu $(HDR)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -DMPCB_PROFILE -c $< -o $@
$(PROF_OUT): build/prof_kernels.o build/mpcb_capi.o build/mpcb_planner_host.o
	$(NVCC) -shared -o $@ $^ -cudart shared -Xlinker -rpath=$(CUDA_HOME)/lib64

# Example mains on top of include/motion_planner.hpp (ports of the reference's examples/*.cpp call sequences)
examples: build/offline_trajectory build/mpc_benchmark
build/offline_trajectory: examples/offline_trajectory.cpp include/motion_planner.hpp $(OUT)
	$(CXX) -O2 -std=c++17 -Wall -Iinclude $< -o $@ -L$(dir $(OUT)) -lmpcb -Wl,-rpath,'$$ORIGIN/../$(dir $(OUT))'
build/mpc_benchmark: examples/benchmark.cpp include/motion_planner.hpp $(OUT)
	$(CXX) -O2 -std=c++17 -Wall -Iinclude $< -o $@ -L$(dir $(OUT)) -lmpcb -Wl,-rpath,'$$ORIGIN/../$(dir $(OUT))'

# The reference's own example mains, compiled UNCHANGED from /root/reference (never copied) against shim/ (a minimal Eigen / pinocchio
# surface + an Eigen-typed MotionPlanner adapter over include/motion_planner.hpp).  Only where the reference tree is mounted.
REF ?= /root/reference
examples-unchanged: build/ref_offline_trajectory build/ref_mpc_benchmark
build/ref_offline_trajectory: $(REF)/examples/offline_trajectory.cpp shim/motionPlanner.hpp shim/Eigen/Dense include/motion_planner.hpp $(OUT)
	$(CXX) -O2 -std=c++17 -Ishim -Iinclude $< -o $@ -L$(dir $(OUT)) -lmpcb -Wl,-rpath,'$$ORIGIN/../$(dir $(OUT))'
build/ref_mpc_benchmark: $(REF)/examples/benchmark.cpp shim/motionPlanner.hpp shim/Eigen/Dense include/motion_planner.hpp $(OUT)
	$(CXX) -O2 -std=c++17 -Ishim -Iinclude $< -o $@ -L$(dir $(OUT)) -lmpcb -Wl,-rpath,'$$ORIGIN/../$(dir $(OUT))'
