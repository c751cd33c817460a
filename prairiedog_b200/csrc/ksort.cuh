// ksort.cuh -- shared constants of the fused occurrence-index sort (ksort.cu).
#pragma once
#include "traverse.cuh"

namespace pdg {

static constexpr int KS_THREADS = 512;
static constexpr int KS_WARPS = KS_THREADS / 32;
static constexpr int KS_ITEMS = 8;
static constexpr int KS_TILE = KS_THREADS * KS_ITEMS;   // 4096 items per CTA
static constexpr int KS_MAX_PASSES = 4;                 // k <= 16: at most 32 key bits
static constexpr int kDirectMaxBits = 26;               // row-start table indexed by code (k <= 13)

struct KsPasses {
    int n;
    int shift[KS_MAX_PASSES];
    uint32_t mask[KS_MAX_PASSES];
};

}  // namespace pdg
