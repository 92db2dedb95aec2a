# Build of the B200-native PCCL: libpccl.so (C++20 core + sm_100a kernels), ccoip_master, C++ tests.
# Everything lands in-tree (pccl_b200/lib, pccl_b200/bin, build/) so that the artefacts travel
# with a snapshot of the repository.
# the image exports CXX=/opt/gcc/bin/g++, a toolchain that links libstdc++ statically; a second copy of
# libstdc++ inside a Python process that already loaded the shared one crashes in iostream code
CXX      := $(shell command -v /usr/bin/g++ || echo g++)
NVCC     ?= /usr/local/cuda/bin/nvcc
CUDA_HOME ?= /usr/local/cuda
BUILD    := build
LIBDIR   := pccl_b200/lib
BINDIR   := pccl_b200/bin

CXXFLAGS := -std=c++20 -O3 -g1 -ffp-contract=off -fPIC -fvisibility=hidden -Wall -Wextra -Wpedantic -Wshadow=local -Iinclude -Icsrc -pthread
NVFLAGS  := -std=c++20 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden \
            -Iinclude -Icsrc --expt-relaxed-constexpr -Xptxas -v
LDFLAGS  := -pthread -Wl,--no-undefined

# make SANITIZE=address|thread NO_CUDA=1 tests   -> instrumented C++ tests under build-<kind>/ (the shipped
# library and the normal build tree are left alone)
ifneq ($(SANITIZE),)
BUILD    := build-$(SANITIZE)
LIBDIR   := $(BUILD)/lib
BINDIR   := $(BUILD)/bin
ifeq ($(SANITIZE),address)
SANFLAGS := -fsanitize=address,undefined -fno-sanitize-recover=undefined
else
SANFLAGS := -fsanitize=$(SANITIZE)
endif
CXXFLAGS := $(subst -O3,-O1,$(CXXFLAGS)) -fno-omit-frame-pointer $(SANFLAGS)
LDFLAGS  := -pthread $(SANFLAGS)
endif

ifeq ($(NO_CUDA),1)
CXXFLAGS += -DPCCL_NO_CUDA
CU_SRCS  :=
CUDA_LIBS :=
else
CXXFLAGS += -I$(CUDA_HOME)/include
CU_SRCS  := $(wildcard csrc/kernels/cuda/*.cu) $(wildcard csrc/device/*.cu)
CUDA_LIBS := -L$(CUDA_HOME)/lib64 -lcudart_static -ldl -lrt
endif

CPP_SRCS := $(wildcard csrc/util/*.cpp csrc/wire/*.cpp csrc/net/*.cpp csrc/topo/*.cpp csrc/master/*.cpp \
              csrc/kernels/cpu/*.cpp csrc/client/*.cpp csrc/device/*.cpp csrc/api/*.cpp)
CPP_OBJS := $(patsubst %.cpp,$(BUILD)/%.o,$(CPP_SRCS))
CU_OBJS  := $(patsubst %.cu,$(BUILD)/%.cu.o,$(CU_SRCS))

TEST_SRCS := $(wildcard tests/cpp/test_*.cpp)
TEST_BINS := $(patsubst tests/cpp/%.cpp,$(BUILD)/tests/%,$(TEST_SRCS))

.PHONY: all lib master tests clean
all: lib master tests
lib: $(LIBDIR)/libpccl.so
master: $(BINDIR)/ccoip_master
tests: $(TEST_BINS) $(BUILD)/tests/c99_compat

$(BUILD)/%.o: %.cpp
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -MMD -MP -c $< -o $@

# nvcc writes no dependency files here: every CUDA object depends on every kernel header
CU_HDRS := $(wildcard csrc/kernels/cuda/*.cuh csrc/kernels/*.hpp include/*.h)
$(BUILD)/%.cu.o: %.cu $(CU_HDRS)
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)

$(LIBDIR)/libpccl.so: $(CPP_OBJS) $(CU_OBJS)
	@mkdir -p $(LIBDIR)
	$(CXX) -shared -o $@ $^ $(LDFLAGS) $(CUDA_LIBS)

$(BINDIR)/ccoip_master: tools/ccoip_master.cpp $(LIBDIR)/libpccl.so
	@mkdir -p $(BINDIR)
	$(CXX) $(CXXFLAGS) -fvisibility=default $< -o $@ -L$(LIBDIR) -lpccl -Wl,-rpath,'$$ORIGIN/../lib' -pthread

# C++ tests link the objects directly so they can reach internal classes
$(BUILD)/tests/%: tests/cpp/%.cpp $(CPP_OBJS) $(CU_OBJS)
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -Itests/cpp $< $(CPP_OBJS) $(CU_OBJS) -o $@ $(LDFLAGS) $(CUDA_LIBS)

# the headers must compile as strict C99
$(BUILD)/tests/c99_compat: tests/cpp/c99_compat.c include/pccl.h include/pccl_kernels.h $(LIBDIR)/libpccl.so
	@mkdir -p $(dir $@)
	gcc -std=c99 -pedantic -Wall -Wextra -Werror -Iinclude $< -o $@ -L$(LIBDIR) -lpccl -Wl,-rpath,'$$ORIGIN/../../$(LIBDIR)'

clean:
	rm -rf $(BUILD) $(LIBDIR)/libpccl.so $(BINDIR)/ccoip_master

-include $(CPP_OBJS:.o=.d)
