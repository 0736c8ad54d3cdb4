# -------------------------
# Makefile.flags.mk
# Turns BACKEND_<config> into CXXFLAGS_<config> / LDFLAGS_<config> / LDLIBS_<config>.
#
# Backends of this tree:
#   b200         the product: host code is plain C++17 over the C ABI in include/dlnb.h,
#                linked against dlnetbench_b200/libdlnb.so (built by its own Makefile with
#                nvcc -gencode arch=compute_100a,code=sm_100a). No external packages.
#   nccl         COMPARISON build (never the product): the same drivers, emulated compute, buffers and timing,
#                but every collective is one NCCL call, as in the reference's CCLCommunicator
#                (cpp/proxy_classes.hpp:135-253). Needs nccl.h + libnccl (2.27.3 in this image) and the
#                CUDA runtime headers; built only where they exist.
#   mpi_cpu_ref  the reference's sources compiled by oracle/Makefile (checker / CPU baseline).
# The reference's other backends (rccl, oneccl, mpi_gpu_cuda, mpi_gpu_hip) are not built here.
# -------------------------

REPO_ROOT  := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))
DLNB_DIR   := $(REPO_ROOT)dlnetbench_b200
DLNB_LIB   := $(DLNB_DIR)/libdlnb.so
NVCC       ?= nvcc
CUDA_HOME  ?= /usr/local/cuda

define backend_flags
CXXFLAGS_$(1) := -O3 -std=c++17 -Wall
LDFLAGS_$(1)  :=
LDLIBS_$(1)   := -lpthread -ldl
ifeq ($$(BACKEND_$(1)),b200)
CXXFLAGS_$(1) += -I$(REPO_ROOT)include -DPROXY_ENABLE_B200
LDFLAGS_$(1)  += -L$(DLNB_DIR) -Wl,-rpath,'$$$$ORIGIN/../../../../dlnetbench_b200' -Wl,-rpath,$(DLNB_DIR)
LDLIBS_$(1)   += -ldlnb
NEEDS_DLNB_$(1) := 1
endif
ifeq ($$(BACKEND_$(1)),nccl)
CXXFLAGS_$(1) += -I$(REPO_ROOT)include -DPROXY_ENABLE_B200 -DPROXY_ENABLE_NCCL -I$(CUDA_HOME)/include
LDFLAGS_$(1)  += -L$(DLNB_DIR) -Wl,-rpath,'$$$$ORIGIN/../../../../dlnetbench_b200' -Wl,-rpath,$(DLNB_DIR)
LDLIBS_$(1)   += -ldlnb -lnccl
NEEDS_DLNB_$(1) := 1
endif
endef

$(foreach cfg,$(CONFIGS) $(MAKECMDGOALS),$(if $(BACKEND_$(cfg)),$(eval $(call backend_flags,$(cfg)))))

$(DLNB_LIB): $(wildcard $(DLNB_DIR)/csrc/*.cu $(DLNB_DIR)/csrc/*.cuh $(DLNB_DIR)/csrc/*.cpp $(DLNB_DIR)/csrc/*.hpp $(DLNB_DIR)/csrc/*.inc) $(REPO_ROOT)include/dlnb.h
	$(MAKE) --no-print-directory -C $(DLNB_DIR)/csrc NVCC=$(NVCC)
