# Builds libedison_b200.so (the product) and the oracle (test infrastructure).
#   make            -> edison_b200/libedison_b200.so   (sm_100a only)
#   make oracle     -> oracle/libedison_oracle.so, and oracle/_ref/*.so when /root/reference exists
#   make emu        -> tests/host_emu/libemu_mfcc.so   (host emulation of the K1 warp program)
NVCC      ?= /usr/local/cuda/bin/nvcc
CC        ?= gcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall -Xptxas -v
CSRC      := edison_b200/csrc
OBJDIR    := build
LIB       := edison_b200/libedison_b200.so

CU_SRCS   := $(CSRC)/mfcc_q15_kernel.cu $(CSRC)/nnom_q7_kernel.cu $(CSRC)/nnom_q7_umma.cu $(CSRC)/edison_api.cu $(CSRC)/stream.cu $(CSRC)/mfcc_float.cu $(CSRC)/cube_f32_kernel.cu $(CSRC)/nnom_generic.cu $(CSRC)/noalloc.cu
HDRS      := $(CSRC)/k2_umma_plan.h $(CSRC)/nnom_blob.h $(CSRC)/q15_arith.cuh $(CSRC)/mfcc_q15_core.cuh $(CSRC)/edison_internal.h include/edison_b200.h
OBJS      := $(OBJDIR)/mfcc_q15_kernel.o $(OBJDIR)/nnom_q7_kernel.o $(OBJDIR)/nnom_q7_umma.o $(OBJDIR)/edison_api.o $(OBJDIR)/stream.o $(OBJDIR)/mfcc_float.o $(OBJDIR)/cube_f32_kernel.o $(OBJDIR)/nnom_generic.o $(OBJDIR)/noalloc.o $(OBJDIR)/tables.o $(OBJDIR)/firmware_api.o

all: $(LIB)

$(OBJDIR):
	mkdir -p $(OBJDIR)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS) | $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)
	@grep -E "registers|spill" $(OBJDIR)/$*.ptxas.log | sed 's/^/  [$*] /' || true

$(OBJDIR)/tables.o: $(CSRC)/tables.cpp $(HDRS) | $(OBJDIR)
	$(CXX) -O2 -std=c++17 -fPIC -ffp-contract=off -fno-strict-aliasing -Wall -I/usr/local/cuda/include -c $< -o $@

$(OBJDIR)/firmware_api.o: $(CSRC)/firmware_api.c include/edison_b200.h | $(OBJDIR)
	$(CC) -O2 -std=c99 -fPIC -Wall -Wextra -c $< -o $@

$(LIB): $(OBJS) $(CSRC)/exports.map
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -Xlinker --version-script=$(CSRC)/exports.map -Xlinker -z,defs -ldl

oracle:
	$(MAKE) -C oracle all
	$(MAKE) -C oracle ref

emu: tests/host_emu/libemu_mfcc.so tests/host_emu/libemu_k2.so
tests/host_emu/libemu_k2.so: tests/host_emu/emu_k2.cpp $(CSRC)/k2_umma_plan.h $(CSRC)/nnom_blob.h
	$(CXX) -O2 -std=c++17 -Wall -Wno-unknown-pragmas -fPIC -shared -I$(CSRC) tests/host_emu/emu_k2.cpp -o $@
tests/host_emu/libemu_mfcc.so: tests/host_emu/emu_mfcc.cpp $(CSRC)/tables.cpp $(HDRS)
	$(CXX) -O2 -std=c++17 -ffp-contract=off -fno-strict-aliasing -fPIC -shared -I/usr/local/cuda/include -I$(CSRC) tests/host_emu/emu_mfcc.cpp $(CSRC)/tables.cpp -o $@

clean:
	rm -rf $(OBJDIR) $(LIB) tests/host_emu/libemu_mfcc.so tests/host_emu/libemu_k2.so

.PHONY: all oracle emu clean
