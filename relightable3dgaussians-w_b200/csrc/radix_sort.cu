// radix_sort.cu - hand-written stable LSD "onesweep" radix sort of (key, u32 value) pairs for sm_100a.
//
// Replaces cub::DeviceRadixSort::SortPairs at the reference's two call sites
//   (tile|depth) keys : diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:303-308
//   Morton codes      : simple-knn/simple_knn.cu:210-213
// Semantics that parity depends on: sorts on bits [0, end_bit) only, and is STABLE (equal keys keep
// their input order = ascending Gaussian index).
//
// Structure (one read of the keys for all histograms, then one read + one write per 8-bit pass):
//   1. radix_histogram_kernel   : per-pass digit histograms (smem atomics, one global flush per block)
//   2. radix_scan_kernel        : exclusive scan of each pass's 256 bins -> global digit bases
//   3. onesweep_pass_kernel x n : tile of 4096 pairs per CTA; warp-level digit match from ballots, ranking (stable),
//                                 per-digit decoupled look-back across tiles (chained scan fused into
//                                 the scatter pass), smem reorder, coalesced scatter.
// HBM-bound: algorithmic bytes = n*(sizeof(K)+ (24 per pass for u64/u32 pairs)).
#include <cstdlib>
#include <type_traits>
#include "common.cuh"

namespace r3dgw {

constexpr int OS_THREADS = 256;
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_ITEMS_MIN = 8;                   // smallest tile any instantiation uses (sizes the look-back workspace)

struct SortPasses {
    int n_pass;
    int shift[SORT_MAX_PASSES];
    uint32_t mask[SORT_MAX_PASSES];
};

static SortPasses make_passes(int end_bit) {
    SortPasses sp;
    sp.n_pass = (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
    if (sp.n_pass < 1) sp.n_pass = 1;
    for (int i = 0; i < sp.n_pass; ++i) {
        sp.shift[i] = i * SORT_RADIX_BITS;
        int bits = end_bit - sp.shift[i];
        if (bits > SORT_RADIX_BITS) bits = SORT_RADIX_BITS;
        if (bits < 0) bits = 0;
        sp.mask[i] = (1u << bits) - 1u;
    }
    return sp;
}

// workspace layout: hist u32[MAX_PASSES][256] | tile counters u32[MAX_PASSES] (+pad) | status u64[n_pass][tiles][256]
static inline size_t os_tiles(int n, int items) { return ((size_t)n + OS_THREADS * items - 1) / (OS_THREADS * items); }
size_t sort_workspace_bytes(int n) {
    size_t head = SORT_MAX_PASSES * SORT_RADIX * sizeof(uint32_t) + 256;
    return align_up(head + os_tiles(n, OS_ITEMS_MIN) * SORT_RADIX * sizeof(unsigned long long) * SORT_MAX_PASSES, 256);
}

template <typename KeyT>
__global__ void __launch_bounds__(256) radix_histogram_kernel(const KeyT* __restrict__ keys, int n,
                                                              SortPasses sp, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
    for (int i = threadIdx.x; i < sp.n_pass * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        KeyT k = keys[i];
#pragma unroll
        for (int p = 0; p < SORT_MAX_PASSES; ++p)
            if (p < sp.n_pass) atomicAdd(&s_hist[p * SORT_RADIX + ((uint32_t)(k >> sp.shift[p]) & sp.mask[p])], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < sp.n_pass * SORT_RADIX; i += blockDim.x) {
        uint32_t v = s_hist[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// one block per pass: in-place exclusive scan of 256 bins
__global__ void __launch_bounds__(SORT_RADIX) radix_scan_kernel(uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_w[SORT_RADIX / 32];
    uint32_t* h = hist + blockIdx.x * SORT_RADIX;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t v = h[threadIdx.x], incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (unsigned w = 0; w < warp; ++w) base += s_w[w];
    h[threadIdx.x] = base + incl - v;
}

__device__ __forceinline__ void os_status_store(unsigned long long* p, unsigned flag, unsigned v) {
    unsigned long long w = ((unsigned long long)flag << 32) | v;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long os_status_load(const unsigned long long* p) {
    unsigned long long w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}

template <typename KeyT, int OS_ITEMS>
struct OsSmem {
    static constexpr int OS_TILE = OS_THREADS * OS_ITEMS;
    uint32_t warp_hist[OS_WARPS][SORT_RADIX];   // per-warp digit counts, then exclusive-across-warps offsets
    uint32_t digit_start[SORT_RADIX];           // tile-local start of each digit in the reordered tile
    uint32_t bin_offset[SORT_RADIX];            // global position = bin_offset[d] + local position
    uint32_t scan_tmp[OS_WARPS];
    uint32_t tile_id;
    KeyT keys[OS_TILE];
    uint32_t vals[OS_TILE];
};

template <typename KeyT, int OS_ITEMS, bool BALLOT_MATCH = false>
__global__ void __launch_bounds__(OS_THREADS) onesweep_pass_kernel(
    const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, KeyT* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, int n, int shift, uint32_t mask,
    const uint32_t* __restrict__ digit_base,     // exclusive-scanned histogram of this pass [256]
    uint32_t* __restrict__ tile_counter,         // dynamic tile id allocator (forward progress)
    unsigned long long* __restrict__ status)     // [tiles][256] look-back words
{
    extern __shared__ __align__(16) unsigned char os_smem_raw[];
    constexpr int OS_TILE = OS_THREADS * OS_ITEMS;
    OsSmem<KeyT, OS_ITEMS>& s = *reinterpret_cast<OsSmem<KeyT, OS_ITEMS>*>(os_smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) s.tile_id = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < OS_WARPS * SORT_RADIX; i += OS_THREADS) (&s.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s.tile_id;
    const int tile_base = (int)tile * OS_TILE;
    const int warp_base = tile_base + (int)warp * (OS_ITEMS * 32);

    // ---- load (warp-striped: item i of lane l sits at warp_base + i*32 + l) --------------------
    KeyT key[OS_ITEMS];
    uint32_t val[OS_ITEMS];
    uint32_t rank[OS_ITEMS];
#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        int g = warp_base + i * 32 + (int)lane;
        if (g < n) {
            key[i] = keys_in[g];
            val[i] = vals_in[g];
        } else {
            key[i] = (KeyT)0;
            val[i] = 0;
        }
    }

    // ---- stable warp-level ranking --------------------------------------------------------------
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        int g = warp_base + i * 32 + (int)lane;
        uint32_t d = (g < n) ? ((uint32_t)(key[i] >> shift) & mask) : SORT_RADIX;   // sentinel for the tail
        uint32_t peers;
        if constexpr (BALLOT_MATCH) {
            // lanes holding the same digit, from 8 + 1 ballots: constant cost, whereas MATCH.ANY iterates over the
            // distinct values in the warp - and the low tile-id bits of consecutive duplicates are all distinct
            // (voting only the pass's significant bits, or skipping the validity vote on full tiles, puts uniform branches
            // into the ranking loop and was measured slower: 0.105 -> 0.114 ms per step)
            const bool valid = d < SORT_RADIX;
            const uint32_t bv = __ballot_sync(0xffffffffu, valid);
            peers = valid ? bv : ~bv;
#pragma unroll
            for (int bit = 0; bit < SORT_RADIX_BITS; ++bit) {
                const bool on = (d >> bit) & 1u;
                const uint32_t bal = __ballot_sync(0xffffffffu, on);
                peers &= on ? bal : ~bal;
            }
        } else {
            peers = __match_any_sync(0xffffffffu, d);
        }
        uint32_t leader = __ffs(peers) - 1;
        uint32_t before = 0;
        if (lane == leader && d < SORT_RADIX)     // one shared-memory atomic instead of a dependent load + store
            before = atomicAdd(&s.warp_hist[warp][d], (uint32_t)__popc(peers));
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[i] = before + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // ---- per digit (thread == digit): cross-warp offsets, tile count, look-back ----------------
    {
        const uint32_t d = tid;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < OS_WARPS; ++w) {
            uint32_t c = s.warp_hist[w][d];
            s.warp_hist[w][d] = run;
            run += c;
        }
        const uint32_t count = run;
        unsigned long long* my = status + (size_t)tile * SORT_RADIX + d;
        if (tile == 0) os_status_store(my, 2u, count);
        else os_status_store(my, 1u, count);

        // block-wide exclusive scan of count over digits -> digit_start
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s.scan_tmp[warp] = incl;
        __syncthreads();
        uint32_t wbase = 0;
        for (unsigned w = 0; w < warp; ++w) wbase += s.scan_tmp[w];
        const uint32_t start = wbase + incl - count;
        s.digit_start[d] = start;

        uint32_t excl = 0;
        if (tile > 0) {
            // look back over the predecessors' (flag, count) words, LB of them in flight per step: the walk is a
            // chain of dependent L2 round trips (~0.5 us each), which bounds a pass over a small array
            // (the P-sized depth sort: 245 tiles) long before bandwidth does
            constexpr int LB = 4;
            int j = (int)tile - 1;
            bool done = false;
            while (!done) {
                unsigned long long w[LB];
#pragma unroll
                for (int u = 0; u < LB; ++u)
                    w[u] = (j - u >= 0) ? os_status_load(status + (size_t)(j - u) * SORT_RADIX + d) : (2ull << 32);
#pragma unroll
                for (int u = 0; u < LB; ++u) {
                    if (done) break;
                    const unsigned flag = (unsigned)(w[u] >> 32);
                    if (flag == 0) break;              // not published yet: reload from this predecessor
                    excl += (uint32_t)w[u];
                    --j;
                    if (flag == 2) done = true;
                }
            }
            os_status_store(my, 2u, excl + count);
        }
        s.bin_offset[d] = digit_base[d] + excl - start;
    }
    __syncthreads();

    // ---- reorder through shared memory ----------------------------------------------------------
#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        int g = warp_base + i * 32 + (int)lane;
        if (g < n) {
            uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            uint32_t pos = s.digit_start[d] + s.warp_hist[warp][d] + rank[i];
            s.keys[pos] = key[i];
            s.vals[pos] = val[i];
        }
    }
    __syncthreads();
    const int valid = min(OS_TILE, n - tile_base);
#pragma unroll 4
    for (int pos = tid; pos < valid; pos += OS_THREADS) {
        KeyT k = s.keys[pos];
        uint32_t d = (uint32_t)(k >> shift) & mask;
        uint32_t g = s.bin_offset[d] + pos;
        keys_out[g] = k;
        vals_out[g] = s.vals[pos];
    }
}

template <typename KeyT>
static int sort_pairs_impl(KeyT* keys_in, KeyT* keys_alt, uint32_t* vals_in, uint32_t* vals_alt, int n,
                           int end_bit, char* workspace, cudaStream_t stream, KeyT** keys_out,
                           uint32_t** vals_out) {
    *keys_out = keys_in;
    *vals_out = vals_in;
    if (n <= 0) return 0;
    SortPasses sp = make_passes(end_bit);
    if (sp.n_pass > SORT_MAX_PASSES) return R3DGW_ERR_UNSUPPORTED;
    // pairs per CTA tile = 256 x items: R3DGW_OS_ITEMS = 8 / 12 / 16 (A/B; 16 is the measured default)
    static const int items = [] { const char* e = getenv("R3DGW_OS_ITEMS"); const int v = e ? atoi(e) : 16;
                                  return (v == 8 || v == 12) ? v : 16; }();
    const size_t tiles = os_tiles(n, items);
    uint32_t* hist = reinterpret_cast<uint32_t*>(workspace);
    uint32_t* counters = hist + SORT_MAX_PASSES * SORT_RADIX;
    unsigned long long* status = reinterpret_cast<unsigned long long*>(
        workspace + SORT_MAX_PASSES * SORT_RADIX * sizeof(uint32_t) + 256);
    const size_t status_per_pass = tiles * SORT_RADIX;
    R3DGW_CUDA_TRY(cudaMemsetAsync(workspace, 0,
        SORT_MAX_PASSES * SORT_RADIX * sizeof(uint32_t) + 256 + status_per_pass * sp.n_pass * sizeof(unsigned long long),
        stream));

    int hb = (int)std::min<size_t>((size_t)(n + 256 * 16 - 1) / (256 * 16), (size_t)NUM_SMS * 8);
    {
        ProfScope ps(g_sort_prof_kid >= 0 ? g_sort_prof_kid : KID_SORT_HIST, stream);
        radix_histogram_kernel<KeyT><<<hb, 256, 0, stream>>>(keys_in, n, sp, hist);
    }
    R3DGW_LAUNCH_CHECK();
    {
        ProfScope ps(g_sort_prof_kid >= 0 ? g_sort_prof_kid : KID_SORT_SCAN, stream);
        radix_scan_kernel<<<sp.n_pass, SORT_RADIX, 0, stream>>>(hist);
    }
    R3DGW_LAUNCH_CHECK();

    // R3DGW_OS_MATCH=any: MATCH.ANY for the in-warp digit match instead of the ballot form (A/B)
    static const bool ballot_match = [] { const char* e = getenv("R3DGW_OS_MATCH"); return !(e && e[0] == 'a'); }();
    auto launch_pass = [&](auto tag, auto btag, int p, const KeyT* kin, const uint32_t* vin, KeyT* kout, uint32_t* vout) -> int {
        constexpr int IT = decltype(tag)::value;
        constexpr bool BM = decltype(btag)::value;
        const int smem = (int)sizeof(OsSmem<KeyT, IT>);
        R3DGW_CUDA_TRY(cudaFuncSetAttribute((onesweep_pass_kernel<KeyT, IT, BM>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        onesweep_pass_kernel<KeyT, IT, BM><<<(unsigned)tiles, OS_THREADS, smem, stream>>>(
            kin, vin, kout, vout, n, sp.shift[p], sp.mask[p], hist + p * SORT_RADIX, counters + p,
            status + (size_t)p * status_per_pass);
        return 0;
    };

    KeyT* kin = keys_in; KeyT* kout = keys_alt;
    uint32_t* vin = vals_in; uint32_t* vout = vals_alt;
    for (int p = 0; p < sp.n_pass; ++p) {
        {
            ProfScope ps(g_sort_prof_kid >= 0 ? g_sort_prof_kid : KID_SORT_PASS, stream);
            int rc;
            if (items == 8) rc = launch_pass(std::integral_constant<int, 8>{}, std::true_type{}, p, kin, vin, kout, vout);
            else if (items == 12) rc = launch_pass(std::integral_constant<int, 12>{}, std::true_type{}, p, kin, vin, kout, vout);
            else if (ballot_match) rc = launch_pass(std::integral_constant<int, 16>{}, std::true_type{}, p, kin, vin, kout, vout);
            else rc = launch_pass(std::integral_constant<int, 16>{}, std::false_type{}, p, kin, vin, kout, vout);
            if (rc) return rc;
        }
        R3DGW_LAUNCH_CHECK();
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    *keys_out = kin;
    *vals_out = vin;
    return 0;
}

int sort_pairs_u64_u32(uint64_t* keys_in, uint64_t* keys_alt, uint32_t* vals_in, uint32_t* vals_alt, int n,
                       int end_bit, char* workspace, cudaStream_t stream, uint64_t** keys_out,
                       uint32_t** vals_out) {
    return sort_pairs_impl<uint64_t>(keys_in, keys_alt, vals_in, vals_alt, n, end_bit, workspace, stream,
                                     keys_out, vals_out);
}
int sort_pairs_u32_u32(uint32_t* keys_in, uint32_t* keys_alt, uint32_t* vals_in, uint32_t* vals_alt, int n,
                       int end_bit, char* workspace, cudaStream_t stream, uint32_t** keys_out,
                       uint32_t** vals_out) {
    return sort_pairs_impl<uint32_t>(keys_in, keys_alt, vals_in, vals_alt, n, end_bit, workspace, stream,
                                     keys_out, vals_out);
}

int sort_num_passes(int end_bit) { return make_passes(end_bit).n_pass; }

}  // namespace r3dgw
