// The handle behind `struct rwkv_b200_matrix *` (include/rwkv_b200.h codec / mat-vec hooks); shared by
// api_ext.cpp and no_alloc.cu.
#pragma once
#include "context.h"

struct rwkv_b200_matrix {
    int device = 0;
    std::vector<rb::DevMat> copies;
    std::vector<void *> allocations;
    rb::Act x; int x_rows = 0;
    float * y = nullptr; size_t y_cap = 0;
    cudaStream_t stream = nullptr;
    size_t next = 0;
    rb::MatmulScratch scratch;              // K-split partial tiles (tensor-core hook)
};

