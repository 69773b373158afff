// Optional per-kernel CUDA-event timing inside the library (used by bench.py for the roofline
// numbers and the launch count).  Off by default; when off the only cost is one branch.
#pragma once
#include <cuda_runtime.h>

void fdgs_prof_begin(const char* name, cudaStream_t s);
void fdgs_prof_end(cudaStream_t s);
void fdgs_count_launch();

struct FdgsProfScope {
    cudaStream_t s;
    FdgsProfScope(const char* name, cudaStream_t st) : s(st) { fdgs_prof_begin(name, st); }
    ~FdgsProfScope() { fdgs_prof_end(s); }
};
#define FDGS_PROF(name, stream) FdgsProfScope prof_scope__(name, stream)

// fdgs_set_sm_limit (include/fdgs.h): upper bound on the CTAs of the persistent one-CTA-per-SM deformation kernels
int fdgs_sm_limit_value();
