# Builds the product: libraytracing_b200.so (host C + CUDA sm_100a, C ABI of include/rt_api.h) and bin/raytracing.
# `make oracle` builds the CPU oracle (test infrastructure), `make ref` the real reference into oracle/_ref/.
NVCC ?= nvcc
CC := gcc
CXX := g++
PKG := cuda_raytracing_b200
DEV := $(PKG)/csrc/device
HOST := $(PKG)/csrc/host
BUILD := build
LIB := $(PKG)/libraytracing_b200.so
GENCODE := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(GENCODE) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Iinclude -I$(DEV)
CFLAGS := -O2 -ffp-contract=off -fPIC -std=c11 -Wall -Iinclude
CXXFLAGS := -O2 -fPIC -std=c++17 -Wall -Iinclude -I$(DEV) -I/usr/local/cuda/include

all: $(LIB) bin/raytracing

$(BUILD):
	mkdir -p $(BUILD) bin

$(BUILD)/rt_kernels.o: $(DEV)/rt_kernels.cu $(DEV)/*.cuh $(DEV)/*.h include/*.h | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(BUILD)/rt_api.o: $(DEV)/rt_api.cu $(DEV)/*.h include/*.h | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(BUILD)/rt_group.o: $(DEV)/rt_group.cu $(DEV)/*.h include/*.h | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(BUILD)/rt_pack.o: $(DEV)/rt_pack.cpp $(DEV)/*.h include/*.h | $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@
$(BUILD)/%.o: $(HOST)/%.c include/*.h | $(BUILD)
	$(CC) $(CFLAGS) -c $< -o $@

OBJS := $(BUILD)/rt_kernels.o $(BUILD)/rt_api.o $(BUILD)/rt_group.o $(BUILD)/rt_pack.o $(BUILD)/rt_world.o $(BUILD)/rt_scenes.o $(BUILD)/rt_host.o

$(LIB): $(OBJS)
	$(NVCC) $(GENCODE) -shared -o $@ $(OBJS) -lm -ldl

bin/raytracing: $(BUILD)/main.o $(LIB)
	$(CC) -o $@ $(BUILD)/main.o -L$(PKG) -lraytracing_b200 -Wl,-rpath,'$$ORIGIN/../$(PKG)' -lm

oracle:
	$(MAKE) -C oracle liboracle.so
ref:
	bash oracle/build_ref.sh

clean:
	rm -rf $(BUILD) $(LIB) bin/raytracing

.PHONY: all oracle ref clean
