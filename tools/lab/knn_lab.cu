// knn_lab.cu — micro-lab for the Hamming kNN inner loop (not shipped; compares instruction-mix variants on the GPU).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o knn_lab knn_lab.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s @%d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint32_t maj(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint32_t mad(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ void ins(uint32_t key, uint32_t &b1, uint32_t &b2) { const uint32_t hi = max(key, b1); b1 = min(key, b1); b2 = min(b2, hi); }
__device__ __forceinline__ void cp16(void *s, const void *g) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((uint32_t)__cvta_generic_to_shared(s)), "l"(g)); }

// POPCV: 8 = plain, 5 = three CSA, 4 = four CSA.  BR: 1 = threshold branch, 0 = unconditional insert.
template <int POPCV, int BR>
__device__ __forceinline__ uint32_t keyof(const uint32_t *q, uint4 ta, uint4 tb, uint32_t tkey)
{
    const uint32_t x0 = q[0] ^ ta.x, x1 = q[1] ^ ta.y, x2 = q[2] ^ ta.z, x3 = q[3] ^ ta.w;
    const uint32_t x4 = q[4] ^ tb.x, x5 = q[5] ^ tb.y, x6 = q[6] ^ tb.z, x7 = q[7] ^ tb.w;
    if (POPCV == 8) {
        uint32_t d = __popc(x0) + __popc(x1) + __popc(x2) + __popc(x3) + __popc(x4) + __popc(x5) + __popc(x6) + __popc(x7);
        return (d << 18) + tkey;
    }
    const uint32_t s0 = xor3(x0, x1, x2), c0 = maj(x0, x1, x2);
    const uint32_t s1 = xor3(x3, x4, x5), c1 = maj(x3, x4, x5);
    const uint32_t s2 = xor3(s0, s1, x6), c2 = maj(s0, s1, x6);
    if (POPCV == 15) { // 5 POPC, weighted sum as an IMAD chain
        uint32_t kk = mad(__popc(s2), 1u << 18, tkey);
        kk = mad(__popc(x7), 1u << 18, kk);
        kk = mad(__popc(c0), 1u << 19, kk);
        kk = mad(__popc(c1), 1u << 19, kk);
        return mad(__popc(c2), 1u << 19, kk);
    }
    if (POPCV == 16) { // 6 POPC, two CSA only
        uint32_t kk = mad(__popc(s0), 1u << 18, tkey);
        kk = mad(__popc(s1), 1u << 18, kk);
        kk = mad(__popc(x6), 1u << 18, kk);
        kk = mad(__popc(x7), 1u << 18, kk);
        kk = mad(__popc(c0), 1u << 19, kk);
        return mad(__popc(c1), 1u << 19, kk);
    }
    if (POPCV == 5) {
        const uint32_t ones = __popc(s2) + __popc(x7), twos = __popc(c0) + __popc(c1) + __popc(c2);
        return ((ones + 2u * twos) << 18) + tkey;
    }
    const uint32_t t = xor3(c0, c1, c2), f = maj(c0, c1, c2);
    uint32_t kk = mad(__popc(s2), 1u << 18, tkey);
    kk = mad(__popc(x7), 1u << 18, kk);
    kk = mad(__popc(t), 1u << 19, kk);
    return mad(__popc(f), 1u << 20, kk);
}

// XOR-linear forms (the shipped compare, kfs_bf.cu): q = {q0, q1, Q012, q3, q4, Q345, Q0..6, q7}, qx = Q0..7; the train row's
// combinations t012, t345, t06, t07 are formed once per row per thread.  carry(a, b, s) = maj(a, b, a ^ b ^ s) = LOP3 0xd4.
__device__ __forceinline__ uint32_t carry(uint32_t a, uint32_t b, uint32_t s) { uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xd4;" : "=r"(r) : "r"(a), "r"(b), "r"(s)); return r; }
__device__ __forceinline__ uint32_t andn(uint32_t a, uint32_t b) { uint32_t r; asm("lop3.b32 %0, %1, %2, %2, 0x30;" : "=r"(r) : "r"(a), "r"(b)); return r; } // a & ~b
// POPCV 25: 11 LOP3 + 5 POPC (shipped).  24: 14 LOP3 + 4 POPC: the two weight-1 words go through a half adder whose sum is XOR-linear
// (Q0..7 ^ T0..7) and whose carry is s2 & ~sum; the three carries through one more full adder.  26: as 24, weighted sum as an IMAD
// chain with the multipliers in registers (k18, k19, k20: opaque, so that ptxas keeps IMAD = FMA pipe instead of LEA / IADD3 = ALU).
template <int POPCV>
__device__ __forceinline__ uint32_t keylin(const uint32_t *q, uint32_t qx, uint4 ta, uint4 tb, uint32_t t012, uint32_t t345, uint32_t t06, uint32_t t07,
                                           uint32_t tkey, uint32_t k18, uint32_t k19, uint32_t k20)
{
    const uint32_t s0 = q[2] ^ t012, s1 = q[5] ^ t345, s2 = q[6] ^ t06;
    const uint32_t x0 = q[0] ^ ta.x, x1 = q[1] ^ ta.y, x3 = q[3] ^ ta.w, x4 = q[4] ^ tb.x;
    const uint32_t c0 = carry(x0, x1, s0), c1 = carry(x3, x4, s1), c2 = carry(s0, s1, s2);
    if (POPCV == 25) {
        const uint32_t x7 = q[7] ^ tb.w;
        const uint32_t ones = __popc(s2) + __popc(x7), twos = __popc(c0) + __popc(c1) + __popc(c2);
        return ((ones + 2u * twos) << 18) + tkey;
    }
    const uint32_t sum1 = qx ^ t07, h = andn(s2, sum1);
    const uint32_t t = xor3(c0, c1, c2), f = maj(c0, c1, c2);
    if (POPCV == 24) {
        const uint32_t twos = __popc(t) + __popc(h);
        return ((__popc(sum1) + 2u * twos + 4u * __popc(f)) << 18) + tkey;
    }
    uint32_t kk = mad(__popc(sum1), k18, tkey);
    kk = mad(__popc(t), k19, kk);
    kk = mad(__popc(h), k19, kk);
    return mad(__popc(f), k20, kk);
}

template <int POPCV, int BR, int QPT, int TSPLIT, int THREADS, int CHUNK, int UNR>
__global__ void __launch_bounds__(THREADS) k_knn(const uint8_t *__restrict__ Q, const uint8_t *__restrict__ T, int nq, int nt, uint2 *__restrict__ out)
{
    constexpr int QWARPS = THREADS / 32 / TSPLIT, QTILE = QWARPS * 32 * QPT, SLICE = CHUNK / TSPLIT;
    extern __shared__ __align__(16) uint4 s_train[]; // [2][CHUNK*2]
    const uint8_t *query = Q + (size_t)blockIdx.y * nq * 32, *train = T + (size_t)blockIdx.y * nt * 32;
    const int q0 = blockIdx.x * QTILE;
    if (q0 >= nq) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, qgroup = warp % QWARPS, tsplit = warp / QWARPS;
    uint32_t q[QPT][8], qx[QPT], b1[QPT], b2[QPT];
#pragma unroll
    for (int r = 0; r < QPT; r++) {
        const int qi = min(q0 + qgroup * (32 * QPT) + r * 32 + lane, nq - 1);
        const uint4 *src = reinterpret_cast<const uint4 *>(query + (size_t)qi * 32);
        const uint4 a = __ldg(src), b = __ldg(src + 1);
        q[r][0] = a.x; q[r][1] = a.y; q[r][2] = a.z; q[r][3] = a.w; q[r][4] = b.x; q[r][5] = b.y; q[r][6] = b.z; q[r][7] = b.w;
        if (POPCV >= 20) {
            q[r][2] = xor3(a.x, a.y, a.z); q[r][5] = xor3(a.w, b.x, b.y); q[r][6] = xor3(q[r][2], q[r][5], b.z);
            qx[r] = q[r][6] ^ b.w;
        }
        b1[r] = b2[r] = 0xffffffffu;
    }
    uint32_t k18 = 1u << 18, k19 = 1u << 19, k20 = 1u << 20;
    asm volatile("" : "+r"(k18), "+r"(k19), "+r"(k20));
    const int nchunks = (nt + CHUNK - 1) / CHUNK;
    auto issue = [&](int chunk, int stage) {
        const int base = chunk * CHUNK, rows = min(CHUNK, nt - base);
        const uint4 *src = reinterpret_cast<const uint4 *>(train + (size_t)base * 32);
        for (int i = tid; i < rows * 2; i += THREADS) cp16(&s_train[stage * CHUNK * 2 + i], src + i);
        asm volatile("cp.async.commit_group;\n" ::);
    };
    if (nchunks > 0) issue(0, 0);
    for (int c = 0; c < nchunks; c++) {
        const int stage = c & 1;
        if (c + 1 < nchunks) { issue(c + 1, stage ^ 1); asm volatile("cp.async.wait_group 1;\n" ::); }
        else asm volatile("cp.async.wait_group 0;\n" ::);
        __syncthreads();
        const int base = c * CHUNK + tsplit * SLICE, rows = min(SLICE, nt - base);
        const uint4 *tp = &s_train[stage * CHUNK * 2 + tsplit * SLICE * 2];
        int jstart = 0;
        if (BR == 2) {
#pragma unroll UNR
            for (; jstart + 1 < rows; jstart += 2) {
                const uint4 ta = tp[2 * jstart], tb = tp[2 * jstart + 1], tc = tp[2 * jstart + 2], td = tp[2 * jstart + 3];
                const uint32_t tkey = (uint32_t)(base + jstart);
                uint32_t a012 = 0, a345 = 0, a06 = 0, a07 = 0, c012 = 0, c345 = 0, c06 = 0, c07 = 0;
                if (POPCV >= 20) {
                    a012 = xor3(ta.x, ta.y, ta.z); a345 = xor3(ta.w, tb.x, tb.y); a06 = xor3(a012, a345, tb.z); a07 = a06 ^ tb.w;
                    c012 = xor3(tc.x, tc.y, tc.z); c345 = xor3(tc.w, td.x, td.y); c06 = xor3(c012, c345, td.z); c07 = c06 ^ td.w;
                }
#pragma unroll
                for (int r = 0; r < QPT; r++) {
                    const uint32_t kA = POPCV >= 20 ? keylin<POPCV>(q[r], qx[r], ta, tb, a012, a345, a06, a07, tkey, k18, k19, k20) : keyof<POPCV, 0>(q[r], ta, tb, tkey);
                    const uint32_t kB = POPCV >= 20 ? keylin<POPCV>(q[r], qx[r], tc, td, c012, c345, c06, c07, tkey + 1, k18, k19, k20) : keyof<POPCV, 0>(q[r], tc, td, tkey + 1);
                    const uint32_t lo = min(kA, kB), hi = max(kA, kB);
                    const uint32_t mx = max(b1[r], lo);
                    b1[r] = min(b1[r], lo);
                    b2[r] = __vimin3_u32(b2[r], mx, hi);
                }
            }
        }
#pragma unroll UNR
        for (int j = jstart; j < rows; j++) {
            const uint4 ta = tp[2 * j], tb = tp[2 * j + 1];
            const uint32_t tkey = (uint32_t)(base + j);
            uint32_t key[QPT];
            bool upd = false;
            uint32_t t012 = 0, t345 = 0, t06 = 0, t07 = 0;
            if (POPCV >= 20) { t012 = xor3(ta.x, ta.y, ta.z); t345 = xor3(ta.w, tb.x, tb.y); t06 = xor3(t012, t345, tb.z); t07 = t06 ^ tb.w; }
#pragma unroll
            for (int r = 0; r < QPT; r++) {
                key[r] = POPCV >= 20 ? keylin<POPCV>(q[r], qx[r], ta, tb, t012, t345, t06, t07, tkey, k18, k19, k20) : keyof<POPCV, BR>(q[r], ta, tb, tkey);
                if (BR == 1) upd |= key[r] < b2[r]; else ins(key[r], b1[r], b2[r]);
            }
            if (BR == 1 && upd) {
#pragma unroll
                for (int r = 0; r < QPT; r++) ins(key[r], b1[r], b2[r]);
            }
        }
        __syncthreads();
    }
    uint32_t *s_m = reinterpret_cast<uint32_t *>(s_train);
#pragma unroll
    for (int r = 0; r < QPT; r++) {
        const int ql = qgroup * (32 * QPT) + r * 32 + lane;
        s_m[(tsplit * QTILE + ql) * 2 + 0] = b1[r]; s_m[(tsplit * QTILE + ql) * 2 + 1] = b2[r];
    }
    __syncthreads();
    for (int ql = tid; ql < QTILE; ql += THREADS) {
        const int qi = q0 + ql;
        if (qi >= nq) continue;
        uint32_t m1 = 0xffffffffu, m2 = 0xffffffffu;
        for (int s = 0; s < TSPLIT; s++) { ins(s_m[(s * QTILE + ql) * 2], m1, m2); ins(s_m[(s * QTILE + ql) * 2 + 1], m1, m2); }
        out[(size_t)blockIdx.y * nq + qi] = make_uint2(m1, m2);
    }
}

static uint8_t *dQ, *dT; static uint2 *dO; static int NQ = 2000, NT = 2000, NP = 64;
template <int POPCV, int BR, int QPT, int TSPLIT, int THREADS, int CHUNK, int UNR>
static void run(const char *name)
{
    constexpr int QTILE = THREADS / 32 / TSPLIT * 32 * QPT;
    auto kern = k_knn<POPCV, BR, QPT, TSPLIT, THREADS, CHUNK, UNR>;
    const size_t smem = 2 * CHUNK * 32;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
    int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem));
    dim3 grid((NQ + QTILE - 1) / QTILE, NP);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) kern<<<grid, THREADS, smem>>>(dQ, dT, NQ, NT, dO);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; i++) kern<<<grid, THREADS, smem>>>(dQ, dT, NQ, NT, dO);
    CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= reps;
    std::vector<uint2> h((size_t)NP * NQ); CK(cudaMemcpy(h.data(), dO, h.size() * 8, cudaMemcpyDeviceToHost));
    uint64_t cs = 0; for (auto &v : h) cs = cs * 1000003ull + v.x * 31ull + v.y;
    printf("%-34s regs %3d occ %d CTA/SM  %8.1f us  %7.1f Gcmp/s  cs %016llx\n", name, fa.numRegs, occ, ms * 1e3, (double)NP * NQ * NT / ms / 1e6, (unsigned long long)cs);
}
#define RUN(...) run<__VA_ARGS__>(#__VA_ARGS__)

int main(int argc, char **argv)
{
    if (argc > 1) NQ = NT = atoi(argv[1]);
    if (argc > 2) NP = atoi(argv[2]);
    std::vector<uint8_t> hq((size_t)NP * NQ * 32), ht((size_t)NP * NT * 32);
    uint64_t s = 88172645463325252ull;
    auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
    for (auto &b : hq) b = (uint8_t)rnd();
    for (size_t i = 0; i < ht.size(); i++) ht[i] = (i / 32) % 3 == 0 ? hq[i % hq.size()] ^ (uint8_t)(rnd() & rnd() & rnd()) : (uint8_t)rnd();
    CK(cudaMalloc(&dQ, hq.size())); CK(cudaMalloc(&dT, ht.size())); CK(cudaMalloc(&dO, (size_t)NP * NQ * 8));
    CK(cudaMemcpy(dQ, hq.data(), hq.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dT, ht.data(), ht.size(), cudaMemcpyHostToDevice));
    //  POPCV, BR, QPT, TSPLIT, THREADS, CHUNK, UNR
    RUN(5, 0, 4, 8, 256, 512, 2);
    RUN(25, 0, 4, 8, 256, 512, 2);
    RUN(24, 0, 4, 8, 256, 512, 2);
    RUN(26, 0, 4, 8, 256, 512, 2);
    RUN(26, 2, 4, 8, 256, 1024, 2);
    RUN(26, 2, 4, 8, 256, 256, 2);
    RUN(26, 2, 4, 8, 256, 512, 1);
    RUN(26, 2, 4, 8, 256, 512, 2);
    RUN(26, 2, 3, 8, 256, 512, 2);
    RUN(26, 0, 4, 8, 256, 512, 1);
    RUN(26, 0, 4, 8, 256, 512, 4);
    RUN(26, 0, 3, 8, 256, 512, 2);
    RUN(26, 0, 6, 8, 256, 512, 1);
    RUN(26, 0, 4, 4, 128, 512, 2);
    RUN(15, 0, 4, 8, 256, 512, 2);
    RUN(16, 0, 4, 8, 256, 512, 2);
    RUN(4, 0, 4, 8, 256, 512, 2);
    RUN(15, 2, 4, 8, 256, 512, 1);
    RUN(16, 2, 4, 8, 256, 512, 1);
    RUN(4, 2, 4, 8, 256, 512, 1);
    RUN(15, 2, 4, 8, 256, 512, 2);
    RUN(15, 2, 2, 8, 256, 512, 2);
    RUN(15, 2, 3, 8, 256, 512, 1);
    RUN(15, 0, 3, 8, 256, 512, 2);
    RUN(15, 0, 2, 8, 256, 512, 4);
    RUN(15, 0, 4, 4, 128, 512, 2);
    RUN(15, 2, 4, 4, 128, 512, 1);
    return 0;
}
