// oracle/compat/ccutils/cuda/cuda_macros.hpp — TEST INFRASTRUCTURE (oracle). Only reached by
// the reference's CUDA builds (proxy_classes.hpp:8,403-405,434); shape follows its HIP twin,
// cpp/tmp_hip_ccutils.hpp:6-19. The mpi_cpu oracle build never includes it.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CCUTILS_CUDA_CHECK(call)                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(1);                                                                               \
        }                                                                                          \
    } while (0)
#define CCUTILS_CUDA_FREE_SAFE(p)                                                                  \
    do { if (p) { cudaFree(p); (p) = nullptr; } } while (0)
