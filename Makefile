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

HOST_HDRS := $(wildcard $(PKG)/host/*.h) include/nvdb_b200.h
HOST_BINS := $(BINDIR)/nvdb_bench $(BINDIR)/nvdb_gt_build $(BINDIR)/nvdb_convert_f16 $(BINDIR)/nvdb_quantize_i8 \
             $(BINDIR)/nvdb_ivfpq_build $(BINDIR)/nvdb_ivf_eval
HOST_CXX  := $(CXX) -std=c++17 -O2 -fopenmp -Iinclude -I$(PKG)/host
HOST_LD   := -L$(LIBDIR) -lnvdb_b200 -Wl,-rpath,'$$ORIGIN/../lib' -lpthread
host: lib $(HOST_BINS)
$(filter-out $(BINDIR)/nvdb_quantize_i8,$(HOST_BINS)): | $(LIBDIR)/libnvdb_b200.so

$(BINDIR)/nvdb_bench: $(PKG)/host/nvdb_bench.cpp $(HOST_HDRS)
	@mkdir -p $(BINDIR)
	$(HOST_CXX) -o $@ $< $(HOST_LD)

$(BINDIR)/nvdb_gt_build: $(PKG)/host/nvdb_gt_build.cpp $(HOST_HDRS)
	@mkdir -p $(BINDIR)
	$(HOST_CXX) -o $@ $< $(HOST_LD)

$(BINDIR)/nvdb_ivfpq_build: $(PKG)/host/nvdb_ivfpq_build.cpp $(HOST_HDRS)
	@mkdir -p $(BINDIR)
	$(HOST_CXX) -o $@ $< $(HOST_LD)

$(BINDIR)/nvdb_ivf_eval: $(PKG)/host/nvdb_ivf_eval.cpp $(HOST_HDRS)
	@mkdir -p $(BINDIR)
	$(HOST_CXX) -o $@ $< $(HOST_LD)

# one source, two names: the tool picks f32->f16 or ->i8 from argv[0] (the reference ships two separate tools)
$(BINDIR)/nvdb_convert_f16: $(PKG)/host/nvdb_convert.cpp $(HOST_HDRS)
	@mkdir -p $(BINDIR)
	$(HOST_CXX) -o $@ $< $(HOST_LD)

$(BINDIR)/nvdb_quantize_i8: $(BINDIR)/nvdb_convert_f16
	cp $< $@

clean:
	rm -rf build $(LIBDIR) $(BINDIR)
