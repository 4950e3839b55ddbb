// kfs_bf.cu — brute-force Hamming k-NN (k <= 2) on sm_100a: the replacement for
//     ptrMatcher->knnMatch(keyframe->matDescriptors, dpCurrImg, knnMatches, 2)
//         /root/reference/src/lib/VisualOdometry.cpp:39   (BFMatcher(NORM_HAMMING) at :18, ratio test at :46)
// Contract (OpenCV batchDistance + insertion top-K): distance = popcount of the
// 256-bit XOR; per query the K smallest in ascending order, ties -> lower trainIdx first; rows hold
// min(K, Nt) entries; Nt < 2^18.
//
// This is integer work: XOR / LOP3 on the ALU pipe + POPC on the XU pipe, no tensor cores.  Query descriptors live in
// registers (QPT per thread), train descriptors are streamed through shared memory in double-buffered
// chunks (cp.async), every lane of a warp reads the same train row (broadcast LDS.128), and the running
// top-2 is kept as two packed keys  dist << 18 | trainIdx  so that a plain unsigned min reproduces
// "lower index wins ties" exactly.  A CTA splits the train chunk over TSPLIT warp groups and merges the
// partial top-2 lists through shared memory; the merged result is written as cv::DMatch records.
#include "kfs_common.cuh"
#include <float.h>
#include <string.h>
#include <stdlib.h>
#include <vector>
#include <unordered_map>
#include <new>
#include <mutex>

constexpr int BF_THREADS = 256;
constexpr int BF_QPT = 4;                       // queries per thread
constexpr int BF_TSPLIT = 8;                    // warp groups scanning disjoint train slices
constexpr int BF_QWARPS = BF_THREADS / 32 / BF_TSPLIT;
constexpr int BF_QTILE = BF_QWARPS * 32 * BF_QPT; // 128 queries per CTA
constexpr int BF_CHUNK = 512;                   // train rows per smem stage (16 KB)
constexpr int BF_SLICE = BF_CHUNK / BF_TSPLIT;  // rows one warp group scans per stage
constexpr uint32_t BF_INF = 0xffffffffu;
constexpr int BF_MAX_SPLIT = 16;               // train-row splits of one (pair, query tile) across CTAs
constexpr int BF_GRAN = 128;                   // a split is a whole number of 128-row granules (16 rows per warp group)
static_assert(BF_QTILE <= BF_THREADS, "one thread per query in the merge epilogue");


__device__ __forceinline__ void bf_cp_async16(void *smem, const void *gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void bf_cp_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void bf_cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ uint32_t bf_xor3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t bf_maj(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// carry of a full adder from two addends and the SUM: maj(a, b, a ^ b ^ s)
__device__ __forceinline__ uint32_t bf_carry(uint32_t a, uint32_t b, uint32_t s)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xd4;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
    return r;
}

__device__ __forceinline__ uint32_t bf_andn(uint32_t a, uint32_t b) // a & ~b
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %2, 0x30;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t bf_mad(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

__device__ __forceinline__ void bf_top2_insert(uint32_t key, uint32_t &b1, uint32_t &b2)
{
    const uint32_t hi = max(key, b1);
    b1 = min(key, b1);
    b2 = min(b2, hi);
}

// grid (ceil(max_nq / BF_QTILE), npairs, S).  S > 1 splits the TRAIN rows of every (pair, query tile) over S CTAs so that
// small batches still fill the 148 SMs and large ones end in many short waves instead of a few long ones.  The S partial
// top-2 lists of a tile meet in one of two ways:
//   CLUSTER  the S CTAs are one thread-block cluster (1, 1, S): every CTA stores its list straight into the shared memory of
//            the cluster's rank 0 (mapa + st.shared::cluster), one cluster barrier, rank 0 merges and writes the rows — no
//            global scratch, no fence, no atomic, no read-back (the launch sizes the dynamic shared memory: S x 1 KB)
//   !CLUSTER the lists go through `scratch`, and the CTA that arrives last at the tile's counter merges them (the counter is
//            left at zero for the next launch); used when S = 1 and as the fallback when a cluster launch is refused
extern __shared__ __align__(16) uint2 s_bf_part[]; // CLUSTER, rank 0: [S][BF_QTILE] partial lists
template <bool CLUSTER>
__global__ void __launch_bounds__(BF_THREADS, 3)
k_hamming_knn2(const BfPairDev *__restrict__ pairs, int k, uint2 *__restrict__ scratch, int *__restrict__ counters)
{
    __shared__ int s_last;
    __shared__ __align__(16) uint4 s_train[2][BF_CHUNK * 2]; // 2 stages x 512 rows x 32 B
    const BfPairDev P = pairs[blockIdx.y];
    int nq = P.nq, nt = P.nt;
    if (P.d_nq) nq = min(nq, *P.d_nq);
    if (P.d_nt) nt = min(nt, *P.d_nt);
    const int q0 = blockIdx.x * BF_QTILE;
    if (q0 >= nq) return; // uniform over a cluster: its CTAs differ in blockIdx.z only
    // A CTA may touch another CTA's shared memory only once that CTA is known to have started; being launched as one cluster is
    // not that guarantee.  Every CTA announces itself here and waits for the others just before the remote store below, so a CTA
    // with few or no train rows cannot run ahead of rank 0's start, and nobody waits while there is scanning to do.
    if (CLUSTER) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int qgroup = warp % BF_QWARPS, tsplit = warp / BF_QWARPS;

    // query descriptors -> registers (clamped index for the ragged tail; results of clamped rows are dropped).
    // Kept per query: words 0, 1, 3, 4 and the four XOR-linear combinations the carry-save sums need (see the loop).
    uint32_t q[BF_QPT][8];
    uint32_t b1[BF_QPT], b2[BF_QPT];
#pragma unroll
    for (int r = 0; r < BF_QPT; r++) {
        const int qi = min(q0 + qgroup * (32 * BF_QPT) + r * 32 + lane, nq - 1);
        const uint4 *src = reinterpret_cast<const uint4 *>(P.query + (size_t)qi * 32);
        const uint4 a = __ldg(src), b = __ldg(src + 1);
        q[r][0] = a.x; q[r][1] = a.y; q[r][3] = a.w; q[r][4] = b.x;
        q[r][2] = bf_xor3(a.x, a.y, a.z);            // Q012
        q[r][5] = bf_xor3(a.w, b.x, b.y);            // Q345
        q[r][6] = bf_xor3(q[r][2], q[r][5], b.z);    // Q0..6
        q[r][7] = q[r][6] ^ b.w;                     // Q0..7
        b1[r] = BF_INF; b2[r] = BF_INF;
    }

    uint32_t k18 = 1u << 18, k19 = 1u << 19, k20 = 1u << 20; // key = dist << 18 | trainIdx, built by IMADs (see the loop)
    asm volatile("" : "+r"(k18), "+r"(k19), "+r"(k20));
    // this CTA's train rows [t0, t1)
    const int S = gridDim.z, split = blockIdx.z;
    const int seg = S > 1 ? ((nt + BF_GRAN - 1) / BF_GRAN + S - 1) / S * BF_GRAN : nt; // whole granules per split
    const int t0 = min(split * seg, nt), t1 = min(t0 + seg, nt);
    const int nchunks = (t1 - t0 + BF_CHUNK - 1) / BF_CHUNK;
    auto issue = [&](int chunk, int stage) {
        const int base = t0 + chunk * BF_CHUNK;
        const int rows = min(BF_CHUNK, t1 - base);
        const uint4 *src = reinterpret_cast<const uint4 *>(P.train + (size_t)base * 32);
        for (int i = tid; i < rows * 2; i += BF_THREADS) bf_cp_async16(&s_train[stage][i], src + i);
        bf_cp_commit();
    };
    if (nchunks > 0) issue(0, 0);
    for (int c = 0; c < nchunks; c++) {
        const int stage = c & 1;
        if (c + 1 < nchunks) { issue(c + 1, stage ^ 1); bf_cp_wait<1>(); }
        else bf_cp_wait<0>();
        __syncthreads();
        // the rows of this stage are dealt evenly to the 8 warp groups (a short last stage keeps every group busy)
        const int crows = min(BF_CHUNK, t1 - (t0 + c * BF_CHUNK));
        const int slice = (crows + BF_TSPLIT - 1) / BF_TSPLIT;
        const int base = t0 + c * BF_CHUNK + tsplit * slice;
        const int rows = min(slice, t1 - base); // may be <= 0 for the last groups
        const uint4 *tp = &s_train[stage][tsplit * slice * 2];
        // 256-bit Hamming distance with a carry-save (Harley-Seal) front end: the 8 XOR words x_i = q_i ^ t_i are compressed into bit
        // planes so that only FOUR POPC are needed (the XU pipe, 16 lanes per clock and SM, is the kernel's bound):
        //   x0+x1+x2 = s0 + 2 c0;  x3+x4+x5 = s1 + 2 c1;  s0+s1+x6 = s2 + 2 c2;  s2+x7 = u + 2 h;  c0+c1+c2 = t + 2 f
        //   popc(all) = popc(u) + 2 (popc(t) + popc(h)) + 4 popc(f)
        // The SUM words are XOR-linear, so each is ONE xor of a precombined query word and a precombined train word
        // (s0 = Q012 ^ T012, s1 = Q345 ^ T345, s2 = Q0..6 ^ T0..6, u = Q0..7 ^ T0..7), a CARRY of a full adder is a 3-input function of
        // two addends and the sum (the third addend is their xor): c = maj(a, b, a^b^s) = LOP3 0xd4, and the half adder's carry is
        // h = s2 & x7 = s2 & ~u.  x2, x5, x6, x7 are never formed: 14 LOP3 + 4 POPC per compare.
        // The weighted sum is an IMAD chain whose multipliers sit in registers, so that it stays on the FMA pipe (with immediates
        // ptxas turns it into LEA / IADD3 on the ALU pipe, which the LOP3s already fill).  Round 1's form (11 LOP3 + 5 POPC) ran at
        // 783 Gcmp/s in tools/lab/knn_lab.cu, this one at 811, and at 833 with the two-row top-2 update below.
        // The four train-side combinations do not depend on the query: the warp that will scan a slice first rewrites its rows IN PLACE
        // in shared memory, {t0 t1 t2 t3 | t4 t5 t6 t7} -> {t0 t1 T012 t3 | t4 T345 T0..6 T0..7} (t2, t5, t6, t7 enter the compare only
        // through the combinations), one or two rows per lane, so that the scan below forms nothing per row.
        static_assert(BF_QWARPS == 1, "a train slice is scanned by exactly one warp (the in-place rewrite is warp-local)");
        {
            uint4 *tw = &s_train[stage][tsplit * slice * 2];
            for (int rr = lane; rr < rows; rr += 32) {
                uint4 a = tw[2 * rr], b = tw[2 * rr + 1];
                const uint32_t t012 = bf_xor3(a.x, a.y, a.z), t345 = bf_xor3(a.w, b.x, b.y), t06 = bf_xor3(t012, t345, b.z);
                a.z = t012; b.y = t345; b.z = t06; b.w = t06 ^ b.w;
                tw[2 * rr] = a; tw[2 * rr + 1] = b;
            }
            __syncwarp();
        }
        auto key_of = [&](const uint32_t (&qr)[8], const uint4 &ta, const uint4 &tb, uint32_t tkey) -> uint32_t {
            const uint32_t s0 = qr[2] ^ ta.z, s1 = qr[5] ^ tb.y, s2 = qr[6] ^ tb.z, u = qr[7] ^ tb.w;
            const uint32_t x0 = qr[0] ^ ta.x, x1 = qr[1] ^ ta.y, x3 = qr[3] ^ ta.w, x4 = qr[4] ^ tb.x;
            const uint32_t c0 = bf_carry(x0, x1, s0), c1 = bf_carry(x3, x4, s1), c2 = bf_carry(s0, s1, s2);
            const uint32_t h = bf_andn(s2, u), t = bf_xor3(c0, c1, c2), f = bf_maj(c0, c1, c2);
            uint32_t key = bf_mad(__popc(u), k18, tkey);
            key = bf_mad(__popc(t), k19, key);
            key = bf_mad(__popc(h), k19, key);
            return bf_mad(__popc(f), k20, key);
        };
        int j = 0;
        // two train rows per step: the running (b1 <= b2) and the sorted pair (lo <= hi) of the two new keys merge in 5 min / max
        // operations — b1' = min(b1, lo), b2' = min3(b2, max(b1, lo), hi) — instead of 2 x 3
#pragma unroll 2
        for (; j + 1 < rows; j += 2) {
            const uint4 ta = tp[2 * j], tb = tp[2 * j + 1], tc = tp[2 * j + 2], td = tp[2 * j + 3];
            const uint32_t tkey = (uint32_t)(base + j);
#pragma unroll
            for (int r = 0; r < BF_QPT; r++) {
                const uint32_t kA = key_of(q[r], ta, tb, tkey);
                const uint32_t kB = key_of(q[r], tc, td, tkey + 1u);
                const uint32_t lo = min(kA, kB), hi = max(kA, kB);
                const uint32_t mx = max(b1[r], lo);
                b1[r] = min(b1[r], lo);
                b2[r] = __vimin3_u32(b2[r], mx, hi);
            }
        }
        if (j < rows) { // odd row count
            const uint4 ta = tp[2 * j], tb = tp[2 * j + 1];
#pragma unroll
            for (int r = 0; r < BF_QPT; r++) bf_top2_insert(key_of(q[r], ta, tb, (uint32_t)(base + j)), b1[r], b2[r]);
        }
        __syncthreads(); // everyone done with this stage before it is refilled two iterations later
    }

    // merge the TSPLIT partial lists through shared memory (reuses the train buffer)
    uint32_t *s_m = reinterpret_cast<uint32_t *>(&s_train[0][0]); // [TSPLIT][QTILE][2]
#pragma unroll
    for (int r = 0; r < BF_QPT; r++) {
        const int ql = qgroup * (32 * BF_QPT) + r * 32 + lane;
        s_m[(tsplit * BF_QTILE + ql) * 2 + 0] = b1[r];
        s_m[(tsplit * BF_QTILE + ql) * 2 + 1] = b2[r];
    }
    __syncthreads();
    uint32_t m1 = BF_INF, m2 = BF_INF; // thread ql < BF_QTILE: merged top-2 of query q0 + ql over this CTA's rows
    if (tid < BF_QTILE) {
#pragma unroll
        for (int s = 0; s < BF_TSPLIT; s++) {
            bf_top2_insert(s_m[(s * BF_QTILE + tid) * 2 + 0], m1, m2);
            bf_top2_insert(s_m[(s * BF_QTILE + tid) * 2 + 1], m1, m2);
        }
    }
    if (CLUSTER) {
        // blockIdx.z is the rank inside the (1, 1, S) cluster.  Rank 0's s_bf_part is written by nobody but this exchange, so the
        // faster CTAs do not wait for it to finish its own rows; the barrier (release / acquire) publishes the remote stores.
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory"); // pairs with the arrive at kernel entry: rank 0 is running
        if (tid < BF_QTILE) {
            const uint32_t mine = (uint32_t)__cvta_generic_to_shared(&s_bf_part[split * BF_QTILE + tid]);
            uint32_t dst;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(dst) : "r"(mine), "r"(0));
            asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(dst), "r"(m1), "r"(m2) : "memory");
        }
        asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (split != 0) return;
        if (tid < BF_QTILE) {
            m1 = m2 = BF_INF;
            for (int s = 0; s < S; s++) {
                const uint2 v = s_bf_part[s * BF_QTILE + tid];
                bf_top2_insert(v.x, m1, m2);
                bf_top2_insert(v.y, m1, m2);
            }
        }
    } else if (S > 1) {
        const int tile = blockIdx.y * gridDim.x + blockIdx.x;
        uint2 *part = scratch + (size_t)tile * S * BF_QTILE;
        if (tid < BF_QTILE) __stcg(part + split * BF_QTILE + tid, make_uint2(m1, m2));
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = atomicAdd(counters + tile, 1) == S - 1;
        __syncthreads();
        if (!s_last) return;
        if (tid == 0) counters[tile] = 0;
        __threadfence();
        if (tid < BF_QTILE) {
            m1 = m2 = BF_INF;
            for (int s = 0; s < S; s++) {
                const uint2 v = __ldcg(part + s * BF_QTILE + tid);
                bf_top2_insert(v.x, m1, m2);
                bf_top2_insert(v.y, m1, m2);
            }
        }
    }
    const int qi = q0 + tid;
    if (tid < BF_QTILE && qi < nq) {
        kfs_dmatch r0, r1;
        r0.queryIdx = qi; r1.queryIdx = qi;
        int cnt = 0;
        if (m1 != BF_INF) { r0.trainIdx = (int)(m1 & 0x3ffffu); r0.imgIdx = 0; r0.distance = (float)(m1 >> 18); cnt = 1; }
        else { r0.trainIdx = -1; r0.imgIdx = -1; r0.distance = FLT_MAX; }
        if (m2 != BF_INF && k > 1) { r1.trainIdx = (int)(m2 & 0x3ffffu); r1.imgIdx = 0; r1.distance = (float)(m2 >> 18); cnt = 2; }
        else { r1.trainIdx = -1; r1.imgIdx = -1; r1.distance = FLT_MAX; }
        uint4 *o = reinterpret_cast<uint4 *>(P.out + (size_t)qi * k);
        o[0] = *reinterpret_cast<uint4 *>(&r0);
        if (k > 1) o[1] = *reinterpret_cast<uint4 *>(&r1);
        if (P.counts) P.counts[qi] = cnt;
    }
}

// Fused ratio test (VisualOdometry.cpp:46): one CTA per pair (blockIdx.x), ordered compaction of rows passing
// d0 < ratio*d1 (both present).  A single block scan keeps query order, like the reference's loop.
__global__ void __launch_bounds__(1024)
k_ratio_compact(const BfPairDev *__restrict__ pairs, float ratio, kfs_dmatch *__restrict__ good_base, long long good_stride,
                int *__restrict__ n_good)
{
    __shared__ int s_warp[32];
    __shared__ int s_base, s_total;
    const BfPairDev P = pairs[blockIdx.x];
    int nq = P.nq;
    if (P.d_nq) nq = min(nq, *P.d_nq);
    const kfs_dmatch *knn = P.out;
    kfs_dmatch *good = good_base + (long long)blockIdx.x * good_stride;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int q0 = 0; q0 < nq; q0 += 1024) {
        const int q = q0 + tid;
        bool keep = false;
        kfs_dmatch m0;
        m0.queryIdx = 0; m0.trainIdx = -1; m0.imgIdx = -1; m0.distance = 0.f;
        if (q < nq) {
            m0 = knn[2 * q];
            const kfs_dmatch m1 = knn[2 * q + 1];
            keep = m0.trainIdx >= 0 && m1.trainIdx >= 0 && m0.distance < __fmul_rn(ratio, m1.distance);
        }
        const uint32_t b = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(b);
        __syncthreads();
        if (warp == 0) {
            const int v = s_warp[lane];
            int incl = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            s_warp[lane] = incl - v; // exclusive offset of each warp
            if (lane == 31) s_total = incl;
        }
        __syncthreads();
        const int base = s_base + s_warp[warp];
        if (keep) good[base + __popc(b & ((1u << lane) - 1u))] = m0;
        __syncthreads();
        if (tid == 0) s_base += s_total;
        __syncthreads();
    }
    if (tid == 0) n_good[blockIdx.x] = s_base;
}

// ---------------------------------------------------------------------------------------------------
struct kfs_bf {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    BfPairDev *d_pairs = nullptr;
    int pairs_cap = 0;
    std::vector<BfPairDev> h_pairs_last;
    BfPairDev *h_stage = nullptr; // pinned
    uint8_t *d_buf = nullptr;     // staging for the host-buffer entry points
    size_t d_buf_bytes = 0;
    int launches = 0;
    int sm_count = 0;
    uint2 *d_scratch = nullptr;   // [tiles][S][BF_QTILE] partial top-2 lists of split launches
    int *d_counters = nullptr;    // [tiles] arrival counters (zero between launches)
    size_t scratch_tiles = 0, scratch_elems = 0;
    int force_split = 0;          // KFS_BF_SPLIT (tests / experiments): 0 = automatic
    int cluster = 1;              // KFS_BF_CLUSTER: 0 (or a refused cluster launch) = splits merge through global scratch, 1 = clusters for
                                  // single-wave launches, 2 = clusters for every split launch
    int cluster_max = 8;          // largest cluster the device schedules for this kernel (16 once the non-portable size is allowed)
    const kfs_orb *attached = nullptr; // kfs_bf_attach_orb: host rows this extractor just delivered are matched on the device
    int aliased = 0;              // descriptor matrices of the last host-buffer call that were found on the device
};

// Number of train splits.  A batch with >= 4 query tiles per SM is throughput-bound as it is (S = 1).  Smaller launches
// (one 2000 x 2000 pair is 16 tiles on 148 SMs) split the train rows over up to BF_MAX_SPLIT CTAs per tile, in whole
// 128-row granules (the kernel deals the rows of a stage evenly to its 8 warp groups).
static int bf_choose_split(const kfs_bf *b, int tiles, int max_nt)
{
    const int nchunks = kfs_div_up(max_nt > 0 ? max_nt : 1, BF_GRAN); // granules
    int s;
    if (b->force_split > 0) s = b->force_split;
    else {
        const int want = 4 * (b->sm_count > 0 ? b->sm_count : 148);
        s = tiles >= want ? 1 : kfs_div_up(want, tiles);
    }
    if (s > BF_MAX_SPLIT) s = BF_MAX_SPLIT;
    if (s > nchunks) s = nchunks;
    if (s < 1) s = 1;
    return kfs_div_up(nchunks, kfs_div_up(nchunks, s)); // drop splits that would get no chunk
}

static kfs_status bf_reserve_split(kfs_bf *b, size_t tiles, int S)
{
    if (S <= 1) return KFS_OK;
    const size_t elems = tiles * (size_t)S * BF_QTILE;
    if (tiles > b->scratch_tiles) {
        KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
        if (b->d_counters) cudaFree(b->d_counters);
        b->d_counters = nullptr; b->scratch_tiles = 0;
        const size_t cap = tiles + tiles / 2 + 64;
        KFS_CUDA_TRY(cudaMalloc(&b->d_counters, cap * sizeof(int)));
        KFS_CUDA_TRY(cudaMemset(b->d_counters, 0, cap * sizeof(int)));
        b->scratch_tiles = cap;
    }
    if (elems > b->scratch_elems) {
        KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
        if (b->d_scratch) cudaFree(b->d_scratch);
        b->d_scratch = nullptr; b->scratch_elems = 0;
        const size_t cap = elems + elems / 2;
        KFS_CUDA_TRY(cudaMalloc(&b->d_scratch, cap * sizeof(uint2)));
        b->scratch_elems = cap;
    }
    return KFS_OK;
}

static kfs_status bf_launch_kernel(kfs_bf *b, const BfPairDev *d_table, int npairs, int max_nq, int max_nt, int k, cudaStream_t st)
{
    const int qtiles = kfs_div_up(max_nq, BF_QTILE);
    const int S = bf_choose_split(b, qtiles * npairs, max_nt);
    kfs_status s = bf_reserve_split(b, (size_t)qtiles * npairs, S);
    if (s) return s;
    cudaError_t e = cudaErrorUnknown;
    // a cluster's CTAs are scheduled together on one GPC: right for a launch that is a single wave (one pair, 2000 x 2000: 16 tiles x
    // 16 splits = 256 CTAs, 21.6 instead of 22.5 us), wrong for one that runs in several waves (config 4's one frame, 10 pairs x 16 tiles x
    // 4 splits = 640 CTAs: 81 instead of 73 us measured with clusters)
    const bool one_wave = (long long)qtiles * npairs * S <= 3LL * (b->sm_count > 0 ? b->sm_count : 148);
    if (S > 1 && b->cluster && S <= b->cluster_max && (one_wave || b->cluster == 2)) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(qtiles, npairs, S); cfg.blockDim = dim3(BF_THREADS); cfg.stream = st;
        cfg.dynamicSmemBytes = (size_t)S * BF_QTILE * sizeof(uint2);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = (unsigned)S;
        cfg.attrs = at; cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, k_hamming_knn2<true>, d_table, k, b->d_scratch, b->d_counters);
        if (e != cudaSuccess) { cudaGetLastError(); b->cluster = 0; } // refused (size / co-scheduling): global merge from now on
    }
    if (e != cudaSuccess) {
        k_hamming_knn2<false><<<dim3(qtiles, npairs, S), BF_THREADS, 0, st>>>(d_table, k, b->d_scratch, b->d_counters);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_hamming_knn2 launch: %s", cudaGetErrorString(e));
    b->launches++;
    return KFS_OK;
}

static kfs_status bf_reserve_pairs(kfs_bf *b, int n)
{
    if (n <= b->pairs_cap) return KFS_OK;
    if (b->d_pairs) { KFS_CUDA_TRY(cudaStreamSynchronize(b->stream)); cudaFree(b->d_pairs); b->d_pairs = nullptr; }
    if (b->h_stage) { cudaFreeHost(b->h_stage); b->h_stage = nullptr; }
    const int cap = n < 64 ? 64 : n * 2;
    KFS_CUDA_TRY(cudaMalloc(&b->d_pairs, sizeof(BfPairDev) * (size_t)cap));
    KFS_CUDA_TRY(cudaMallocHost(&b->h_stage, sizeof(BfPairDev) * (size_t)cap));
    b->pairs_cap = cap;
    b->h_pairs_last.clear();
    return KFS_OK;
}

static kfs_status bf_reserve_buf(kfs_bf *b, size_t bytes)
{
    if (bytes <= b->d_buf_bytes) return KFS_OK;
    if (b->d_buf) { KFS_CUDA_TRY(cudaStreamSynchronize(b->stream)); cudaFree(b->d_buf); b->d_buf = nullptr; }
    const size_t cap = kfs_align_up(bytes + bytes / 2, 1 << 20);
    KFS_CUDA_TRY(cudaMalloc(&b->d_buf, cap));
    b->d_buf_bytes = cap;
    return KFS_OK;
}

static kfs_status bf_check_pairs(const kfs_bf_pair *pairs, int npairs, int k, bool device)
{
    if (!pairs || npairs < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "pairs is NULL or npairs < 0");
    if (k != 1 && k != 2) KFS_FAIL(KFS_ERR_UNSUPPORTED, "knnMatch: only k = 1 or 2 is built (the reference uses k = 2)");
    if (npairs > 65535) KFS_FAIL(KFS_ERR_UNSUPPORTED, "more than 65535 pairs in one call");
    for (int i = 0; i < npairs; i++) {
        const kfs_bf_pair &p = pairs[i];
        if (p.nq < 0 || p.nt < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: negative row count", i);
        if (p.nt >= (1 << 18)) KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: train rows %d >= 2^18 (OpenCV's IMGIDX_ONE limit)", i, p.nt);
        if (p.nq > 0 && (!p.query || !p.out)) KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: query/out is NULL", i);
        if (p.nq > 0 && p.nt > 0 && !p.train) KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: train is NULL", i);
        if (!device && (p.d_nq || p.d_nt)) KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: d_nq/d_nt are device-only", i);
        if (device && (((uintptr_t)p.query | (uintptr_t)p.train | (uintptr_t)p.out) & 15))
            KFS_FAIL(KFS_ERR_INVALID_ARG, "pair %d: device pointers must be 16-byte aligned", i);
    }
    return KFS_OK;
}

static kfs_status bf_launch(kfs_bf *b, const BfPairDev *h, int npairs, int k)
{
    int max_nq = 0, max_nt = 0;
    for (int i = 0; i < npairs; i++) {
        max_nq = h[i].nq > max_nq ? h[i].nq : max_nq;
        max_nt = h[i].nt > max_nt ? h[i].nt : max_nt;
    }
    if (npairs == 0 || max_nq == 0) return KFS_OK;
    kfs_status s = bf_reserve_pairs(b, npairs);
    if (s) return s;
    const bool same = (int)b->h_pairs_last.size() == npairs && memcmp(b->h_pairs_last.data(), h, sizeof(BfPairDev) * (size_t)npairs) == 0;
    if (!same) {
        KFS_CUDA_TRY(cudaStreamSynchronize(b->stream)); // pinned staging may still be in flight
        memcpy(b->h_stage, h, sizeof(BfPairDev) * (size_t)npairs);
        KFS_CUDA_TRY(cudaMemcpyAsync(b->d_pairs, b->h_stage, sizeof(BfPairDev) * (size_t)npairs, cudaMemcpyHostToDevice, b->stream));
        b->h_pairs_last.assign(h, h + npairs);
    }
    return bf_launch_kernel(b, b->d_pairs, npairs, max_nq, max_nt, k, b->stream);
}

// ---- library-internal hooks for kfs_replay.cu (not exported) ------------------------------------------
// kNN over a pair table that already lives in device memory, on an explicit stream.
kfs_status kfs_bf_launch_table(kfs_bf *b, const BfPairDev *d_table, int npairs, int max_nq, int k, cudaStream_t st)
{
    if (npairs == 0 || max_nq == 0) return KFS_OK;
    return bf_launch_kernel(b, d_table, npairs, max_nq, max_nq, k, st); // replay: query and train share the capacity
}
const BfPairDev *kfs_bf_device_table(const kfs_bf *b) { return b->d_pairs; }
// Ratio test of every pair of the table: good_base[pair * good_stride ..], n_good[pair].
kfs_status kfs_bf_launch_ratio_table(kfs_bf *b, const BfPairDev *d_table, int npairs, float ratio, kfs_dmatch *good_base, long long good_stride,
                                     int *n_good, cudaStream_t st)
{
    if (npairs == 0) return KFS_OK;
    k_ratio_compact<<<npairs, 1024, 0, st>>>(d_table, ratio, good_base, good_stride, n_good);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_ratio_compact launch: %s", cudaGetErrorString(e));
    b->launches++;
    return KFS_OK;
}

extern "C" {

kfs_status kfs_bf_create(int norm_type, int cross_check, int device, kfs_bf **out)
{
    if (!out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_create: out is NULL");
    *out = nullptr;
    if (norm_type != KFS_NORM_HAMMING) KFS_FAIL(KFS_ERR_UNSUPPORTED, "only NORM_HAMMING (6) is built (VisualOdometry.cpp:18)");
    if (cross_check) KFS_FAIL(KFS_ERR_UNSUPPORTED, "crossCheck=true is not built (the reference uses the default false)");
    const int ndev = kfs_device_count();
    if (ndev == 0) KFS_FAIL(KFS_ERR_NO_DEVICE, "no CUDA device visible; this library has no CPU path");
    if (device < 0 || device >= ndev) KFS_FAIL(KFS_ERR_INVALID_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    kfs_bf *b = new (std::nothrow) kfs_bf();
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "out of host memory");
    b->device = device;
    KfsDeviceGuard guard(device);
    cudaError_t e = guard.ok ? cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking) : cudaErrorInvalidDevice;
    if (e != cudaSuccess) { delete b; KFS_FAIL(KFS_ERR_CUDA, "kfs_bf_create on device %d: %s", device, cudaGetErrorString(e)); }
    cudaDeviceGetAttribute(&b->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (const char *fs = getenv("KFS_BF_SPLIT")) b->force_split = atoi(fs);
    if (const char *fc = getenv("KFS_BF_CLUSTER")) b->cluster = atoi(fc); // 0 never, 1 single-wave launches (default), 2 every split launch
    if (b->cluster) {
        // static 32 KB + up to 16 KB of partial lists: above the 48 KB default, and 16 CTAs per cluster is a non-portable size
        const cudaError_t e1 = cudaFuncSetAttribute(k_hamming_knn2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BF_MAX_SPLIT * BF_QTILE * (int)sizeof(uint2));
        const cudaError_t e2 = cudaFuncSetAttribute(k_hamming_knn2<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e1 != cudaSuccess) { b->cluster = 0; cudaGetLastError(); }
        b->cluster_max = e2 == cudaSuccess ? BF_MAX_SPLIT : 8;
        if (e2 != cudaSuccess) cudaGetLastError();
    }
    b->stream = b->own_stream;
    *out = b;
    return KFS_OK;
}

void kfs_bf_destroy(kfs_bf *b)
{
    if (!b) return;
    KfsDeviceGuard guard(b->device);
    cudaStreamSynchronize(b->stream);
    if (b->d_pairs) cudaFree(b->d_pairs);
    if (b->h_stage) cudaFreeHost(b->h_stage);
    if (b->d_buf) cudaFree(b->d_buf);
    if (b->d_scratch) cudaFree(b->d_scratch);
    if (b->d_counters) cudaFree(b->d_counters);
    if (b->own_stream) cudaStreamDestroy(b->own_stream);
    delete b;
}

kfs_status kfs_bf_set_stream(kfs_bf *b, void *s)
{
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "bf handle is NULL");
    KFS_DEVICE_GUARD(b->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    b->stream = s ? (cudaStream_t)s : b->own_stream;
    return KFS_OK;
}
void *kfs_bf_get_stream(kfs_bf *b) { return b ? (void *)b->stream : nullptr; }
kfs_status kfs_bf_sync(kfs_bf *b)
{
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "bf handle is NULL");
    KFS_DEVICE_GUARD(b->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return KFS_OK;
}
int kfs_bf_last_launches(const kfs_bf *b) { return b ? b->launches : 0; }

kfs_status kfs_bf_attach_orb(kfs_bf *b, kfs_orb *orb)
{
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "bf handle is NULL");
    b->attached = orb;
    return KFS_OK;
}
int kfs_bf_last_aliased(const kfs_bf *b) { return b ? b->aliased : 0; }

kfs_status kfs_bf_knn_match_pairs_device(kfs_bf *b, const kfs_bf_pair *pairs, int npairs, int k)
{
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "bf handle is NULL");
    kfs_status s = bf_check_pairs(pairs, npairs, k, true);
    if (s) return s;
    KFS_DEVICE_GUARD(b->device);
    b->launches = 0;
    std::vector<BfPairDev> h((size_t)npairs);
    for (int i = 0; i < npairs; i++) {
        h[i].query = pairs[i].query; h[i].train = pairs[i].train; h[i].d_nq = pairs[i].d_nq; h[i].d_nt = pairs[i].d_nt;
        h[i].out = pairs[i].out; h[i].counts = pairs[i].counts; h[i].nq = pairs[i].nq; h[i].nt = pairs[i].nt;
    }
    return bf_launch(b, h.data(), npairs, k);
}

// Host-buffer entry point.  Every distinct host descriptor matrix (keyed by pointer + row count; a keyframe that
// is the query of several pairs, or a frame that is train of one pair and query of another, is uploaded once)
// gets a fixed-size device slot; runs of equally sized matrices at a constant host stride (the rows of one
// [frames][capacity][32] array, the usual layout) travel as ONE 2-D copy instead of one copy per pair, and the
// results come back the same way.
struct BfBufKey {
    const uint8_t *p; int rows;
    bool operator==(const BfBufKey &o) const { return p == o.p && rows == o.rows; }
};
struct BfBufKeyHash {
    size_t operator()(const BfBufKey &k) const { return (size_t)(((uint64_t)(uintptr_t)k.p * 0x9E3779B97F4A7C15ull) ^ (uint64_t)(uint32_t)k.rows); }
};
struct BfCopyRun { const uint8_t *h; size_t d_off; size_t bytes; size_t hstride; int count; };

static void bf_add_to_runs(std::vector<BfCopyRun> &runs, const uint8_t *h, size_t d_off, size_t bytes, size_t dstride)
{
    if (!runs.empty()) {
        BfCopyRun &r = runs.back();
        if (r.bytes == bytes && d_off == r.d_off + (size_t)r.count * dstride) {
            if (r.count == 1 && h > r.h && (size_t)(h - r.h) >= bytes) { r.hstride = (size_t)(h - r.h); r.count = 2; return; }
            if (r.count > 1 && h == r.h + (size_t)r.count * r.hstride) { r.count++; return; }
        }
    }
    runs.push_back(BfCopyRun{h, d_off, bytes, bytes, 1});
}

kfs_status kfs_bf_knn_match_pairs(kfs_bf *b, const kfs_bf_pair *pairs, int npairs, int k)
{
    if (!b) KFS_FAIL(KFS_ERR_INVALID_ARG, "bf handle is NULL");
    kfs_status s = bf_check_pairs(pairs, npairs, k, false);
    if (s) return s;
    if (npairs == 0) return KFS_OK;
    KFS_DEVICE_GUARD(b->device);
    b->launches = 0;
    int max_rows = 0, max_nq = 0;
    for (int i = 0; i < npairs; i++) {
        max_rows = pairs[i].nq > max_rows ? pairs[i].nq : max_rows;
        max_rows = pairs[i].nt > max_rows ? pairs[i].nt : max_rows;
        max_nq = pairs[i].nq > max_nq ? pairs[i].nq : max_nq;
    }
    const size_t in_slot = kfs_align_up((size_t)max_rows * 32 + 1, 256);
    const size_t out_slot = kfs_align_up((size_t)max_nq * k * sizeof(kfs_dmatch) + 1, 256);
    const size_t cnt_slot = kfs_align_up((size_t)max_nq * sizeof(int) + 1, 256);
    std::unordered_map<BfBufKey, int, BfBufKeyHash> slot_of; // (pointer, rows) -> input slot
    slot_of.reserve((size_t)npairs * 2);
    std::vector<BfCopyRun> up, down, down_cnt;
    std::vector<int> sq((size_t)npairs), st((size_t)npairs);
    int nslots = 0;
    std::vector<const uint8_t *> alias; // per slot: where the rows already are on the device (attached extractor), or nullptr
    b->aliased = 0;
    auto slot_for = [&](const uint8_t *h, int rows) {
        const BfBufKey key{h, rows};
        auto it = slot_of.find(key);
        if (it != slot_of.end()) return it->second;
        const int sl = nslots++;
        slot_of.emplace(key, sl);
        const uint8_t *dev = b->attached ? kfs_orb_alias_desc(b->attached, h, rows, b->device) : nullptr;
        alias.push_back(dev);
        if (dev) b->aliased++;
        else if (rows > 0) bf_add_to_runs(up, h, (size_t)sl * in_slot, (size_t)rows * 32, in_slot);
        return sl;
    };
    for (int i = 0; i < npairs; i++) { sq[i] = slot_for(pairs[i].query, pairs[i].nq); st[i] = slot_for(pairs[i].train, pairs[i].nt); }
    const size_t o_out = (size_t)nslots * in_slot, o_cnt = o_out + (size_t)npairs * out_slot;
    s = bf_reserve_buf(b, o_cnt + (size_t)npairs * cnt_slot + 256);
    if (s) return s;
    for (const BfCopyRun &r : up) {
        if (r.count == 1) KFS_CUDA_TRY(cudaMemcpyAsync(b->d_buf + r.d_off, r.h, r.bytes, cudaMemcpyHostToDevice, b->stream));
        else KFS_CUDA_TRY(cudaMemcpy2DAsync(b->d_buf + r.d_off, in_slot, r.h, r.hstride, r.bytes, (size_t)r.count, cudaMemcpyHostToDevice, b->stream));
    }
    std::vector<BfPairDev> h((size_t)npairs);
    for (int i = 0; i < npairs; i++) {
        h[i].query = alias[(size_t)sq[i]] ? alias[(size_t)sq[i]] : b->d_buf + (size_t)sq[i] * in_slot;
        h[i].train = alias[(size_t)st[i]] ? alias[(size_t)st[i]] : b->d_buf + (size_t)st[i] * in_slot;
        h[i].d_nq = nullptr; h[i].d_nt = nullptr;
        h[i].out = (kfs_dmatch *)(b->d_buf + o_out + (size_t)i * out_slot); h[i].counts = (int *)(b->d_buf + o_cnt + (size_t)i * cnt_slot);
        h[i].nq = pairs[i].nq; h[i].nt = pairs[i].nt;
        if (pairs[i].nq) {
            bf_add_to_runs(down, (const uint8_t *)pairs[i].out, o_out + (size_t)i * out_slot, (size_t)pairs[i].nq * k * sizeof(kfs_dmatch), out_slot);
            if (pairs[i].counts) bf_add_to_runs(down_cnt, (const uint8_t *)pairs[i].counts, o_cnt + (size_t)i * cnt_slot, (size_t)pairs[i].nq * sizeof(int), cnt_slot);
        }
    }
    s = bf_launch(b, h.data(), npairs, k);
    if (s) return s;
    for (int pass = 0; pass < 2; pass++) {
        const std::vector<BfCopyRun> &runs = pass ? down_cnt : down;
        const size_t dstride = pass ? cnt_slot : out_slot;
        for (const BfCopyRun &r : runs) {
            if (r.count == 1) KFS_CUDA_TRY(cudaMemcpyAsync((void *)r.h, b->d_buf + r.d_off, r.bytes, cudaMemcpyDeviceToHost, b->stream));
            else KFS_CUDA_TRY(cudaMemcpy2DAsync((void *)r.h, r.hstride, b->d_buf + r.d_off, dstride, r.bytes, (size_t)r.count, cudaMemcpyDeviceToHost, b->stream));
        }
    }
    KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return KFS_OK;
}

kfs_status kfs_bf_knn_match(kfs_bf *b, const uint8_t *query, int nq, const uint8_t *train, int nt, int k, kfs_dmatch *out, int *counts)
{
    kfs_bf_pair p;
    p.query = query; p.nq = nq; p.d_nq = nullptr; p.train = train; p.nt = nt; p.d_nt = nullptr; p.out = out; p.counts = counts;
    return kfs_bf_knn_match_pairs(b, &p, 1, k);
}

kfs_status kfs_bf_knn_match_ratio(kfs_bf *b, const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio, kfs_dmatch *good, int *n_good)
{
    if (!b || !n_good) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_ratio: bad arguments");
    *n_good = 0;
    kfs_bf_pair p;
    p.query = query; p.nq = nq; p.d_nq = nullptr; p.train = train; p.nt = nt; p.d_nt = nullptr; p.out = good; p.counts = nullptr;
    kfs_status s = bf_check_pairs(&p, 1, 2, false);
    if (s) return s;
    if (nq == 0) return KFS_OK;
    KFS_DEVICE_GUARD(b->device);
    b->launches = 0;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t r = off; off = kfs_align_up(off + bytes, 256); return r; };
    const size_t oq = take((size_t)nq * 32), ot = take((size_t)nt * 32), oo = take((size_t)nq * 2 * sizeof(kfs_dmatch));
    const size_t og = take((size_t)nq * sizeof(kfs_dmatch)), on = take(2 * sizeof(int));
    s = bf_reserve_buf(b, off + 256);
    if (s) return s;
    KFS_CUDA_TRY(cudaMemcpyAsync(b->d_buf + oq, query, (size_t)nq * 32, cudaMemcpyHostToDevice, b->stream));
    if (nt) KFS_CUDA_TRY(cudaMemcpyAsync(b->d_buf + ot, train, (size_t)nt * 32, cudaMemcpyHostToDevice, b->stream));
    BfPairDev h;
    h.query = b->d_buf + oq; h.train = b->d_buf + ot; h.d_nq = nullptr; h.d_nt = nullptr;
    h.out = (kfs_dmatch *)(b->d_buf + oo); h.counts = nullptr; h.nq = nq; h.nt = nt;
    s = bf_launch(b, &h, 1, 2);
    if (s) return s;
    k_ratio_compact<<<1, 1024, 0, b->stream>>>(b->d_pairs, ratio, (kfs_dmatch *)(b->d_buf + og), 0, (int *)(b->d_buf + on));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_ratio_compact launch: %s", cudaGetErrorString(e));
    b->launches++;
    int ng[2] = {0, 0};
    KFS_CUDA_TRY(cudaMemcpyAsync(ng, b->d_buf + on, sizeof(ng), cudaMemcpyDeviceToHost, b->stream));
    KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    *n_good = ng[0];
    if (ng[0] > 0) {
        KFS_CUDA_TRY(cudaMemcpyAsync(good, b->d_buf + og, (size_t)ng[0] * sizeof(kfs_dmatch), cudaMemcpyDeviceToHost, b->stream));
        KFS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    }
    return KFS_OK;
}

} // extern "C"

// ---- matching without any allocation -----------------------------------------------------------------
// kfs_bf_knn_match_device_static: the split-train launch of k_hamming_knn2<false> (the variant whose partial top-2 lists meet
// in global scratch) for a caller that has device memory of its own but may not ask the driver for more.  Nothing of the
// reference corresponds to this entry point; see include/kfslam_frontend.h for who needs it and why.
//
// Everything the launch needs beyond the caller's descriptors and result rows lives in ONE file-scope device array:
//   [0, BF_ST_OFF_COUNTERS)                 partial lists, [tile][split][BF_QTILE] uint2           (2 MiB)
//   [BF_ST_OFF_COUNTERS, BF_ST_OFF_PAIR)    one arrival counter per query tile                      (512 B)
//   [BF_ST_OFF_PAIR, BF_ST_BYTES)           the one-entry pair table the kernel reads its pointers from
// The array is part of the module image: each GPU that loads the library carries these ~2 MiB until the process exits.
constexpr int BF_ST_MAX_NQ = KFS_BF_STATIC_MAX_QUERIES;
constexpr int BF_ST_TILES = BF_ST_MAX_NQ / BF_QTILE;
constexpr size_t BF_ST_OFF_COUNTERS = (size_t)BF_ST_TILES * BF_MAX_SPLIT * BF_QTILE * sizeof(uint2);
constexpr size_t BF_ST_OFF_PAIR = BF_ST_OFF_COUNTERS + (size_t)BF_ST_TILES * sizeof(int);
constexpr size_t BF_ST_BYTES = BF_ST_OFF_PAIR + 256;
static_assert(BF_ST_MAX_NQ % BF_QTILE == 0, "the query limit is a whole number of tiles");
static_assert(BF_ST_OFF_COUNTERS % 256 == 0 && BF_ST_OFF_PAIR % 256 == 0 && sizeof(BfPairDev) <= 256, "sections stay 256-byte aligned");
static __device__ __align__(256) unsigned char g_bf_static_scratch[BF_ST_BYTES];

// Stream-ordered preparation of a call, in front of the matching kernel: writes the pair table entry from kernel arguments
// (no host-to-device copy, so no staging buffer either) and zeroes the counters this call's tiles will use.  The matching
// kernel leaves its counters at zero, but the array outlives every call and every caller: a launch that was torn down
// half-way (device reset of another context user, a sticky error) must not poison the next one.  Device code may name the
// array directly; only the host needs cudaGetSymbolAddress.
__global__ void __launch_bounds__(128) k_bf_static_prepare(BfPairDev p, int tiles)
{
    if (threadIdx.x == 0) *reinterpret_cast<BfPairDev *>(g_bf_static_scratch + BF_ST_OFF_PAIR) = p;
    int *c = reinterpret_cast<int *>(g_bf_static_scratch + BF_ST_OFF_COUNTERS);
    for (int i = threadIdx.x; i < tiles; i += blockDim.x) c[i] = 0;
}

// Per-device facts the entry point needs, resolved on the first call for that device (with that device current).  In host
// code `g_bf_static_scratch` is only a handle for the runtime's symbol table, not an address a kernel could use.
constexpr int BF_ST_MAX_DEVICES = 64;
struct BfStaticDevice { unsigned char *base; int sm_count; };
static BfStaticDevice g_bf_static_dev[BF_ST_MAX_DEVICES]; // zero-initialised; base != nullptr = resolved
static std::mutex g_bf_static_mu;

static kfs_status bf_static_device(int device, BfStaticDevice *out)
{
    std::lock_guard<std::mutex> lk(g_bf_static_mu);
    BfStaticDevice &d = g_bf_static_dev[device];
    if (!d.base) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_bf_static_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            KFS_FAIL(KFS_ERR_CUDA, "kfs_bf_knn_match_device_static: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
        }
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) { cudaGetLastError(); KFS_FAIL(KFS_ERR_CUDA, "kfs_bf_knn_match_device_static: SM count of device %d: %s", device, cudaGetErrorString(e)); }
        d.sm_count = sms;
        d.base = static_cast<unsigned char *>(p);
    }
    *out = d;
    return KFS_OK;
}

extern "C" kfs_status kfs_bf_knn_match_device_static(int device, void *cuda_stream, const uint8_t *d_query, int nq, const uint8_t *d_train, int nt,
                                                     int k, kfs_dmatch *d_out, int *d_counts)
{
    if (k != 1 && k != 2) KFS_FAIL(KFS_ERR_UNSUPPORTED, "knnMatch: only k = 1 or 2 is built (the reference uses k = 2)");
    if (nq < 0 || nt < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: negative row count");
    if (nq > BF_ST_MAX_NQ)
        KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: %d query rows exceed KFS_BF_STATIC_MAX_QUERIES = %d, the size the built-in "
                                      "scratch was compiled for; split the query rows over several calls or use kfs_bf_knn_match_pairs_device", nq, BF_ST_MAX_NQ);
    if (nt >= (1 << 18)) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: train rows %d >= 2^18 (OpenCV's IMGIDX_ONE limit)", nt);
    if (nq > 0 && (!d_query || !d_out)) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: query/out is NULL");
    if (nq > 0 && nt > 0 && !d_train) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: train is NULL");
    if (((uintptr_t)d_query | (uintptr_t)d_train | (uintptr_t)d_out) & 15)
        KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_bf_knn_match_device_static: device pointers must be 16-byte aligned");
    const int ndev = kfs_device_count();
    if (ndev == 0) KFS_FAIL(KFS_ERR_NO_DEVICE, "no CUDA device visible; this library has no CPU path");
    if (device < 0 || device >= ndev || device >= BF_ST_MAX_DEVICES) KFS_FAIL(KFS_ERR_INVALID_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    if (nq == 0) return KFS_OK;
    KFS_DEVICE_GUARD(device);
    BfStaticDevice dev;
    kfs_status s = bf_static_device(device, &dev);
    if (s) return s;

    // same split rule as the handle's launches (bf_choose_split), without a handle: a single pair never has enough query tiles
    // to fill the SMs on its own, so its train rows are dealt to S CTAs per tile in whole 128-row granules
    const int tiles = kfs_div_up(nq, BF_QTILE);
    const int nchunks = kfs_div_up(nt > 0 ? nt : 1, BF_GRAN);
    const int want = 4 * (dev.sm_count > 0 ? dev.sm_count : 148);
    int S = tiles >= want ? 1 : kfs_div_up(want, tiles);
    if (S > BF_MAX_SPLIT) S = BF_MAX_SPLIT;
    if (S > nchunks) S = nchunks;
    if (S < 1) S = 1;
    S = kfs_div_up(nchunks, kfs_div_up(nchunks, S));

    BfPairDev p;
    p.query = d_query; p.train = d_train; p.d_nq = nullptr; p.d_nt = nullptr; p.out = d_out; p.counts = d_counts; p.nq = nq; p.nt = nt;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    k_bf_static_prepare<<<1, 128, 0, st>>>(p, tiles);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_bf_static_prepare launch: %s", cudaGetErrorString(e));
    k_hamming_knn2<false><<<dim3(tiles, 1, S), BF_THREADS, 0, st>>>(reinterpret_cast<const BfPairDev *>(dev.base + BF_ST_OFF_PAIR), k,
                                                                     reinterpret_cast<uint2 *>(dev.base),
                                                                     reinterpret_cast<int *>(dev.base + BF_ST_OFF_COUNTERS));
    e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_hamming_knn2 (static scratch) launch: %s", cudaGetErrorString(e));
    return KFS_OK;
}
