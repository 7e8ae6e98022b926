// Shared helpers for the native library: error reporting and CUDA call checking.
#pragma once
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#define BDDL_ERR_INVALID (-1)
#define BDDL_ERR_UNSUPPORTED (-2)
#define BDDL_ERR_TIMEOUT (-3)
#define BDDL_ERR_PARTNER (-4)
#define BDDL_ERR_STOPPED (-5)

// Thread-local message buffer read back through bddl_last_error().
char* bddl_err_buf();
int bddl_fail(int code, const char* fmt, ...);

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define BDDL_CUDA(call)                                                                        \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return bddl_fail((int)e__, "%s:%d %s -> %s", __FILE__, __LINE__, #call,            \
                             cudaGetErrorString(e__));                                         \
    } while (0)
#endif
