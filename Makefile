# Top-level build: the product library (CUDA, sm_100a only) and the C++ host layer.
NVCC   ?= /usr/local/cuda/bin/nvcc
CXX    := /usr/bin/g++
PKG    := nano-vectordb_b200
CSRC   := $(PKG)/csrc
LIBDIR := $(PKG)/lib
BINDIR := $(PKG)/bin
OBJDIR := build/obj
ARCH   := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -std=c++17 -O3 $(ARCH) -lineinfo -Xcompiler -fPIC -ccbin $(CXX) -Iinclude -I$(CSRC)

CU_SRCS := $(wildcard $(CSRC)/*.cu)
CU_OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU_SRCS))

.PHONY: all lib host clean ptxas
all: lib host

lib: $(LIBDIR)/libnvdb_b200.so

$(OBJDIR)/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) include/nvdb_b200.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)

$(LIBDIR)/libnvdb_b200.so: $(CU_OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -ccbin $(CXX) -o $@ $^

HOST_SRCS := $(wildcard $(PKG)/host/*.cpp)
host: lib $(BINDIR)/nvdb_bench

$(BINDIR)/nvdb_bench: $(HOST_SRCS) $(wildcard $(PKG)/host/*.h) include/nvdb_b200.h
	@mkdir -p $(BINDIR)
	@if [ -n "$(HOST_SRCS)" ]; then \
	  $(CXX) -std=c++17 -O2 -fopenmp -Iinclude -I$(PKG)/host -o $@ $(HOST_SRCS) -L$(LIBDIR) -lnvdb_b200 -Wl,-rpath,'$$ORIGIN/../lib' -lpthread; \
	fi

clean:
	rm -rf build $(LIBDIR) $(BINDIR)
