// pipelab.cu -- instruction-throughput micro-benchmarks behind the partition-core design (not product code).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o pipelab pipelab.cu && ./pipelab
// Every test runs CTAS_PER_SM x 148 CTAs of THREADS threads; each thread executes ITERS x U copies of one
// instruction (or a short mix) on 8 independent registers.  Reported: warp instructions per clock per SM
// (from clock64 inside CTA 0) -- 4.0 is the issue limit, 2.0 a half-rate pipe.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

static const int THREADS = 512, CTAS_PER_SM = 2, ITERS = 512;

#define R8(OP) OP(r0) OP(r1) OP(r2) OP(r3) OP(r4) OP(r5) OP(r6) OP(r7)
#define R32(OP) R8(OP) R8(OP) R8(OP) R8(OP)

#define PROLOG                                                                                     \
    uint32_t r0 = threadIdx.x, r1 = r0 * 3 + 1, r2 = r0 * 5 + 2, r3 = r0 * 7 + 3, r4 = r0 * 11 + 4,   \
             r5 = r0 * 13 + 5, r6 = r0 * 17 + 6, r7 = r0 * 19 + 7;                                  \
    uint32_t a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)];                            \
    (void)a; (void)b;                                                                              \
    __syncthreads();                                                                               \
    const long long t0 = clock64();
#define EPILOG                                                                                     \
    const long long t1 = clock64();                                                                \
    out[blockIdx.x * blockDim.x + threadIdx.x] = r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7;             \
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;

#define KERNEL(name, OP)                                                                           \
    __global__ void __launch_bounds__(THREADS) name(const uint32_t *in, uint32_t *out, long long *cyc, int iters) { \
        PROLOG                                                                                     \
        for (int it = 0; it < iters; ++it) { R32(OP) }                                             \
        EPILOG                                                                                     \
    }

#define OP_LOP3(r) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r) : "r"(a), "r"(b));
#define OP_IADD(r) asm volatile("add.u32 %0, %0, %1;" : "+r"(r) : "r"(a));
#define OP_IMAD(r) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r) : "r"(a), "r"(b));
#define OP_SHF(r) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(r) : "r"(a), "r"(b));
#define OP_SHR(r) asm volatile("shr.u32 %0, %0, 1;" : "+r"(r));
#define OP_POPC(r) asm volatile("popc.b32 %0, %0;" : "+r"(r));
#define OP_PRMT(r) asm volatile("prmt.b32 %0, %0, %1, 0x2103;" : "+r"(r) : "r"(a));
#define OP_SEL(r) asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; selp.u32 %0, %0, %2, p;}" : "+r"(r) : "r"(a), "r"(b));
#define OP_SETP(r) asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; @p add.u32 %0, %0, 1;}" : "+r"(r) : "r"(a));
#define OP_VOTE(r) asm volatile("{.reg .pred p; setp.ne.u32 p, %0, 0; vote.sync.ballot.b32 %0, p, 0xffffffff;}" : "+r"(r));
#define OP_ISETP_ONLY(r) asm volatile("{.reg .pred p; setp.ne.u32 p, %0, 0; selp.u32 %0, 1, 2, p;}" : "+r"(r));
#define OP_MATCH(r) asm volatile("match.any.sync.b32 %0, %0, 0xffffffff;" : "+r"(r));
#define OP_SHFL(r) asm volatile("shfl.sync.idx.b32 %0, %0, %1, 0x1f, 0xffffffff;" : "+r"(r) : "r"(a));
#define OP_REDUX(r) asm volatile("redux.sync.or.b32 %0, %0, 0xffffffff;" : "+r"(r));
// predicated-off LOP3: the predicate is false at run time (a == 0 never holds: in[] is non-zero)
#define OP_PREDOFF(r) asm volatile("{.reg .pred p; setp.eq.u32 p, %1, 0; @p lop3.b32 %0, %0, %1, %2, 0x96;}" : "+r"(r) : "r"(a), "r"(b));
#define OP_LOP3_IMAD(r) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;\n\tmad.lo.u32 %0, %0, %1, %2;" : "+r"(r) : "r"(a), "r"(b));
#define OP_LOP3_IADD(r) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;\n\tadd.u32 %0, %0, %1;" : "+r"(r) : "r"(a), "r"(b));

KERNEL(k_lop3, OP_LOP3)
KERNEL(k_iadd, OP_IADD)
KERNEL(k_imad, OP_IMAD)
KERNEL(k_shf, OP_SHF)
KERNEL(k_shr, OP_SHR)
KERNEL(k_popc, OP_POPC)
KERNEL(k_prmt, OP_PRMT)
KERNEL(k_sel, OP_SEL)
KERNEL(k_setp_padd, OP_SETP)
KERNEL(k_setp_vote, OP_VOTE)
KERNEL(k_setp_selp, OP_ISETP_ONLY)
KERNEL(k_match, OP_MATCH)
KERNEL(k_shfl, OP_SHFL)
KERNEL(k_redux, OP_REDUX)
KERNEL(k_setp_predoff_lop3, OP_PREDOFF)
KERNEL(k_lop3_imad, OP_LOP3_IMAD)
KERNEL(k_lop3_iadd, OP_LOP3_IADD)

// ---- one ranking round in isolation: digit in r -> peers mask; NB ballots ----------------------------
// current product form: R2P + NB x (VOTE + 2 predicated LOP3) + combine
template <int NB>
__device__ __forceinline__ uint32_t peers_pred(uint32_t d) {
    uint32_t same = 0xffffffffu, other = 0u;
#pragma unroll
    for (int bb = 0; bb < NB; ++bb)
        asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 m, t;\n\t"
                     "and.b32 t, %2, %3;\n\tsetp.ne.u32 p, t, 0;\n\tvote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
                     "@p and.b32 %0, %0, m;\n\t@!p or.b32 %1, %1, m;\n\t}"
                     : "+r"(same), "+r"(other) : "r"(d), "r"(1u << bb));
    return same & ~other;
}
// xor form: acc |= m ^ sext(bit)  (1 LOP3 + sign extension per bit)
template <int NB>
__device__ __forceinline__ uint32_t peers_xor(uint32_t d) {
    uint32_t acc = 0u;
#pragma unroll
    for (int bb = 0; bb < NB; ++bb) {
        const uint32_t m = __ballot_sync(0xffffffffu, (d >> bb) & 1u);
        const uint32_t sx = (uint32_t)((int32_t)(d << (31 - bb)) >> 31);
        acc |= m ^ sx;
    }
    return ~acc;
}
// select form: t = bit ? m : ~m via one LOP3 with the sign mask: t = m ^ ~sx  -> peers &= t
template <int NB>
__device__ __forceinline__ uint32_t peers_bfe(uint32_t d) {
    uint32_t acc = 0xffffffffu;
#pragma unroll
    for (int bb = 0; bb < NB; ++bb) {
        const uint32_t m = __ballot_sync(0xffffffffu, (d >> bb) & 1u);
        int32_t sx;
        asm("bfe.s32 %0, %1, %2, 1;" : "=r"(sx) : "r"(d), "r"(bb));
        acc &= ~(m ^ (uint32_t)sx);
    }
    return acc;
}
template <int NB>
__device__ __forceinline__ uint32_t peers_match(uint32_t d) { return __match_any_sync(0xffffffffu, d); }

template <int NB, int VAR>
__global__ void __launch_bounds__(THREADS) k_rank(const uint32_t *in, uint32_t *out, long long *cyc, int iters) {
    uint32_t x = in[threadIdx.x & 63] * (threadIdx.x + 1u), acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    const uint32_t lt = (1u << (threadIdx.x & 31)) - 1u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x = x * 1664525u + 1013904223u;
            const uint32_t d = x >> (32 - NB);
            uint32_t p;
            if (VAR == 0) p = peers_pred<NB>(d);
            else if (VAR == 1) p = peers_xor<NB>(d);
            else if (VAR == 2) p = peers_bfe<NB>(d);
            else p = peers_match<NB>(d);
            acc += __popc(p & lt) + (__popc(p) << 8);
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

// ---- shared memory: LDS / STS / ATOMS with random addresses in a per-warp table ----------------------
// MODE 0: LDS.32 random word of a 128-word per-warp table; 1: STS.32 random; 2: LDS.U8 random byte of a 128-byte
// table (32 words); 3: STS.U8 random byte; 4: atomicAdd (no return) random word of a CTA-wide 16K table; 5: atomicAdd
// with return; 6: LDS.32 conflict free; 7: STS.64 random slot of a 2048-slot stage; 8: STS.32 random slot of the stage
template <int MODE>
__global__ void __launch_bounds__(THREADS) k_smem(const uint32_t *in, uint32_t *out, long long *cyc, int iters) {
    extern __shared__ uint32_t sm[];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 16384; i += THREADS) sm[i] = i;
    uint32_t x = in[threadIdx.x & 63] * (threadIdx.x + 1u), acc = 0;
    __syncthreads();
    uint32_t *tab = sm + w * 128;
    unsigned char *tab8 = reinterpret_cast<unsigned char *>(sm) + w * 128;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x = x * 1664525u + 1013904223u;
            const uint32_t d = x >> 25;   // 7 bits
            if (MODE == 0) acc += tab[d];
            else if (MODE == 1) tab[d] = x;
            else if (MODE == 2) acc += tab8[d];
            else if (MODE == 3) tab8[d] = (unsigned char)x;
            else if (MODE == 4) atomicAdd(&sm[x >> 18], 1u);
            else if (MODE == 5) acc += atomicAdd(&sm[x >> 18], 1u);
            else if (MODE == 6) acc += tab[(lane + j) & 127];
            else if (MODE == 7) reinterpret_cast<uint2 *>(sm)[x >> 21] = make_uint2(x, acc);
            else if (MODE == 8) sm[x >> 21] = x;
        }
    }
    const long long t1 = clock64();
    __syncthreads();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + sm[threadIdx.x];
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

// byte stores of different lanes into the same word: do they merge?  Every lane writes byte (lane & 3) of word
// (lane >> 2) of its warp's row, then the row is checked.
__global__ void k_byte_merge(uint32_t *out) {
    __shared__ uint32_t row[8];
    if (threadIdx.x < 8) row[threadIdx.x] = 0;
    __syncwarp();
    reinterpret_cast<unsigned char *>(row)[threadIdx.x] = (unsigned char)(threadIdx.x + 1);
    __syncwarp();
    if (threadIdx.x < 8) out[threadIdx.x] = row[threadIdx.x];
}

typedef void (*kern_t)(const uint32_t *, uint32_t *, long long *, int);

static void run(const char *name, kern_t k, double instr_per_iter, size_t smem, const uint32_t *d_in, uint32_t *d_out,
                long long *d_cyc, int sms) {
    if (smem) CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    k<<<sms * CTAS_PER_SM, THREADS, smem>>>(d_in, d_out, d_cyc, 8);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k<<<sms * CTAS_PER_SM, THREADS, smem>>>(d_in, d_out, d_cyc, ITERS);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    long long cyc;
    CK(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
    const double warps_per_sm = (double)THREADS / 32 * CTAS_PER_SM;
    const double winstr = warps_per_sm * ITERS * instr_per_iter;
    printf("%-28s %9.3f ms  %10lld cyc  %7.3f units/clk/SM  (%.2f clk per unit per SMSP)\n", name, ms, cyc,
           winstr / (double)cyc, (double)cyc * 4.0 / winstr);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main() {
    int dev = 0, sms = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    uint32_t h_in[64];
    for (int i = 0; i < 64; ++i) h_in[i] = 0x9e3779b9u * (i + 1) | 1u;
    uint32_t *d_in, *d_out;
    long long *d_cyc;
    CK(cudaMalloc(&d_in, sizeof(h_in)));
    CK(cudaMalloc(&d_out, (size_t)sms * CTAS_PER_SM * THREADS * 4));
    CK(cudaMalloc(&d_cyc, 8));
    CK(cudaMemcpy(d_in, h_in, sizeof(h_in), cudaMemcpyHostToDevice));
    printf("SMs %d, %d threads x %d CTAs/SM; unit = one warp instruction (or one ranking round / smem access)\n", sms, THREADS,
           CTAS_PER_SM);
#define RUN(k, n) run(#k, k, n, 0, d_in, d_out, d_cyc, sms)
    RUN(k_lop3, 32);
    RUN(k_iadd, 32);
    RUN(k_imad, 32);
    RUN(k_shf, 32);
    RUN(k_shr, 32);
    RUN(k_popc, 32);
    RUN(k_prmt, 32);
    RUN(k_sel, 32);
    RUN(k_setp_padd, 32);
    RUN(k_setp_selp, 32);
    RUN(k_setp_vote, 32);
    RUN(k_setp_predoff_lop3, 32);
    RUN(k_match, 32);
    RUN(k_shfl, 32);
    RUN(k_redux, 32);
    RUN(k_lop3_imad, 32);
    RUN(k_lop3_iadd, 32);
    printf("-- ranking rounds (unit = one round of 32 records) --\n");
    run("rank7 pred (product)", k_rank<7, 0>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank7 xor", k_rank<7, 1>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank7 bfe", k_rank<7, 2>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank7 match.any", k_rank<7, 3>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank8 pred (product)", k_rank<8, 0>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank8 xor", k_rank<8, 1>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank4 pred", k_rank<4, 0>, 8, 0, d_in, d_out, d_cyc, sms);
    run("rank11 pred", k_rank<11, 0>, 8, 0, d_in, d_out, d_cyc, sms);
    printf("-- shared memory (unit = one warp access) --\n");
    const size_t sm_bytes = 16384 * 4;
    run("LDS.32 random (128 words)", k_smem<0>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("STS.32 random (128 words)", k_smem<1>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("LDS.U8 random (128 bytes)", k_smem<2>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("STS.U8 random (128 bytes)", k_smem<3>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("ATOMS.ADD noret random 16K", k_smem<4>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("ATOMS.ADD ret random 16K", k_smem<5>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("LDS.32 conflict free", k_smem<6>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("STS.64 random (2048 slots)", k_smem<7>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    run("STS.32 random (2048 slots)", k_smem<8>, 8, sm_bytes, d_in, d_out, d_cyc, sms);
    k_byte_merge<<<1, 32>>>(d_out);
    CK(cudaDeviceSynchronize());
    uint32_t h[8];
    CK(cudaMemcpy(h, d_out, 32, cudaMemcpyDeviceToHost));
    bool ok = true;
    for (int i = 0; i < 8; ++i) {
        const uint32_t want = (4 * i + 1) | ((4 * i + 2) << 8) | ((4 * i + 3) << 16) | ((uint32_t)(4 * i + 4) << 24);
        if (h[i] != want) ok = false;
    }
    printf("byte stores of different lanes into one word merge: %s\n", ok ? "yes" : "NO");
    return 0;
}
