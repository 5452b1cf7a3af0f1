# Builds the product library (CUDA, sm_100a only) and the CPU oracle (test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
CXX ?= g++
CS := minipath_b200/csrc
LIB := minipath_b200/libminipath_b200.so
ORACLE := oracle/liboracle.so

HOST_FLAGS := -O2 -std=c++17 -fPIC -ffp-contract=off -Wall -Wextra -Wno-unused-parameter
NVCC_FLAGS := -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true \
              -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v

all: $(LIB) $(ORACLE)

$(CS)/host_bvh.o: $(CS)/host_bvh.cpp $(CS)/host_bvh_build.inc $(CS)/host_common.h include/minipath_b200.h
	$(CXX) $(HOST_FLAGS) -c $< -o $@
$(CS)/host_camera.o: $(CS)/host_camera.cpp $(CS)/host_common.h include/minipath_b200.h
	$(CXX) $(HOST_FLAGS) -c $< -o $@
$(CS)/host_render.o: $(CS)/host_render.cpp $(CS)/host_common.h include/minipath_b200.h
	$(CXX) $(HOST_FLAGS) -c $< -o $@
$(CS)/host_image.o: $(CS)/host_image.cpp $(CS)/host_common.h include/minipath_b200.h
	$(CXX) $(HOST_FLAGS) -c $< -o $@
$(CS)/host_group.o: $(CS)/host_group.cpp $(CS)/host_common.h include/minipath_b200.h
	$(CXX) $(HOST_FLAGS) -I/usr/local/cuda/include -c $< -o $@
$(CS)/context.o: $(CS)/context.cu $(CS)/context_render.inc $(CS)/context_probes.inc $(CS)/probes.cuh $(CS)/octet.cuh $(CS)/persist.cuh $(CS)/pool.cuh $(CS)/decode.cuh $(CS)/kernels.cuh $(CS)/wavefront.cuh $(CS)/trace.cuh $(CS)/device_math.cuh include/minipath_b200.h
	$(NVCC) $(NVCC_FLAGS) -c $< -o $@ 2> $(CS)/ptxas.log || (cat $(CS)/ptxas.log; false)

$(CS)/build_device.o: $(CS)/build_device.cu $(CS)/host_common.h include/minipath_b200.h
	$(NVCC) $(NVCC_FLAGS) -c $< -o $@ 2> $(CS)/ptxas_build.log || (cat $(CS)/ptxas_build.log; false)

HOST_OBJS := $(CS)/host_bvh.o $(CS)/host_camera.o $(CS)/host_render.o $(CS)/host_image.o $(CS)/host_group.o
$(LIB): $(HOST_OBJS) $(CS)/context.o $(CS)/build_device.o
	$(NVCC) -shared -o $@ $^ -lpthread -lrt -ldl -cudart static

$(ORACLE): oracle/oracle.cpp oracle/oracle_trace.inc oracle/oracle_render.inc oracle/oracle_build.inc oracle/oracle.h include/minipath_b200.h
	$(CXX) -O2 -std=c++17 -fPIC -shared -mavx2 -mfma -ffp-contract=off -Wall -o $@ oracle/oracle.cpp -lpthread

clean:
	rm -f $(CS)/*.o $(CS)/ptxas.log $(CS)/ptxas_build.log $(LIB) $(ORACLE)
.PHONY: all clean
