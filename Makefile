# Builds the B200-native libpga engine.
#   libpga.a          static archive with relocatable sm_100a device code: the
#                     drop-in for the reference's libpga.a (reference Makefile:11-17;
#                     user programs link "nvcc test.o -L.. -lpga -lcurand")
#   libpga_b200.so    the same objects + demo hooks, device-linked, for C-ABI
#                     bindings (ctypes / cgo / JNI)
# `make oracle` builds the CPU oracle (test infrastructure), `make ref` the
# reference itself and the drop-in demo binaries into oracle/_ref/.
NVCC    ?= nvcc
HOSTCC  := $(shell test -x /usr/bin/gcc && echo /usr/bin/gcc || echo gcc)
ARCH    := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -rdc=true -Xcompiler -fPIC -Iinclude
CFLAGS  := -O2 -std=gnu11 -fPIC -Wall -Wextra -Iinclude
SRC     := libpga_b200/csrc
BUILD   := build

OBJS    := $(BUILD)/pga_host.o $(BUILD)/pga_dev.o
DEMO    := $(BUILD)/pga_demo_hooks.o
KHDRS   := $(SRC)/pga_kernels.cuh $(SRC)/pga_device_utils.cuh $(SRC)/pga_dev.h include/pga.h include/pga_b200.h

all: libpga.a libpga_b200/libpga_b200.so

$(BUILD)/pga_host.o: $(SRC)/pga_host.c $(SRC)/pga_dev.h include/pga.h include/pga_b200.h
	@mkdir -p $(BUILD)
	$(HOSTCC) $(CFLAGS) -c $< -o $@

$(BUILD)/pga_dev.o: $(SRC)/pga_dev.cu $(KHDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -Xptxas -v -dc $< -o $@ 2> $(BUILD)/pga_dev.ptxas.log || (cat $(BUILD)/pga_dev.ptxas.log; false)

$(BUILD)/pga_demo_hooks.o: $(SRC)/pga_demo_hooks.cu include/pga.h include/pga_b200.h
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -dc $< -o $@

libpga.a: $(OBJS)
	rm -f $@
	$(NVCC) $(ARCH) -lib $(OBJS) -o $@

libpga_b200/libpga_b200.so: $(OBJS) $(DEMO)
	$(NVCC) $(ARCH) -shared -Xcompiler -fPIC $(OBJS) $(DEMO) -o $@ -lcudart -ldl

# the same library with every row / score / slot index asserted on the device (-DPGA_DEBUG);
# tests/test_gpu_debug_build.py runs parity cases through it (the pool has no compute-sanitizer)
debug: libpga_b200/libpga_b200_dbg.so

$(BUILD)/pga_dev_dbg.o: $(SRC)/pga_dev.cu $(KHDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -DPGA_DEBUG -dc $< -o $@

libpga_b200/libpga_b200_dbg.so: $(BUILD)/pga_host.o $(BUILD)/pga_dev_dbg.o $(DEMO)
	$(NVCC) $(ARCH) -shared -Xcompiler -fPIC $(BUILD)/pga_host.o $(BUILD)/pga_dev_dbg.o $(DEMO) -o $@ -lcudart -ldl

oracle:
	$(MAKE) -C oracle

ref:
	bash oracle/build_ref.sh

clean:
	rm -rf $(BUILD) libpga.a libpga_b200/libpga_b200.so libpga_b200/libpga_b200_dbg.so

.PHONY: all debug oracle ref clean
