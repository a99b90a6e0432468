// table_kernels.cu -- the data-parallel table passes either side of the HMM hot path, on the device:
//
//   sort / segment      order(cell_id, chr, start) of the long table and the (cell, chromosome) sequence boundaries
//                       (R/callHSCN.R:241-247,266: the sort and the per-cell dplyr::filter scans)            -> sgbk_table_order
//   combineBAFCN reduce [totalcounts > filtern], keyed sum of alleleA / alleleB / totalcounts over
//                       (chr, start, end, cell_id, state, copy), per-cell and per-chromosome bin-count filters
//                       (R/combinedata.R:309-325)                                                           -> sgbk_combine
//   between-pass sums   tarones_Z sums over the modal state's rows (R/callHSCN.R:609-617), infloherror over the A-Hom
//                       rows of the kept cells (R/callHSCN.R:883-902), A x B histogram (fitBB's modal state)     -> sgbk_pass_stats2
//   qc_per_cell         per cell, per allele-specific state with > 10 bins: |B/(A+B) - median(BAF)| averaged with
//                       weights n / N (R/util.R:667-691)                                                   -> sgbk_qc_per_cell
//   fix_assignments     singleton smoothing + per-run allele swap (R/callHSCN.R:1072-1286), as written       -> sgbk_fix_assignments
//
// All of it is HBM-bound integer / byte work: a least-significant-digit radix sort (8 bits per pass, only the passes the
// key range needs), block-wide scans, segmented walks.  Nothing here touches tensor cores.
#include "kernels.h"
#include <cstdio>

#define SGB_FULL 0xffffffffu
static __device__ const double gt_logtab[SGB_LOGTAB_LEN] = SGB_LOGTAB_INIT;
static __device__ const double gt_exptab[SGB_EXPTAB_LEN] = SGB_EXPTAB_INIT;
static __device__ const double gt_lftab[SGB_LFTAB_LEN] = SGB_LFTAB_INIT;

// ------------------------------------------------------------------ exclusive scan of uint32 (three kernels)
#define SCAN_TILE 4096                 // elements per block: 256 threads x 16
__global__ void __launch_bounds__(256) k_scan_tiles(const unsigned* __restrict__ in, unsigned* __restrict__ out, long long n,
                                                    unsigned* __restrict__ tile_sums)
{
    __shared__ unsigned wsum[8];
    const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * 16;
    unsigned v[16], run = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) { v[i] = base + i < n ? in[base + i] : 0u; run += v[i]; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned incl = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(SGB_FULL, incl, d); if (lane >= d) incl += u; }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    unsigned woff = 0;
    for (int w = 0; w < wid; w++) woff += wsum[w];
    unsigned ex = woff + incl - run;
#pragma unroll
    for (int i = 0; i < 16; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
    if (threadIdx.x == 255) tile_sums[blockIdx.x] = ex;
}
__global__ void __launch_bounds__(1024) k_scan_sums(unsigned* sums, int n, unsigned long long* total)
{
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < n; b0 += 1024) {
        const int i = b0 + threadIdx.x;
        const unsigned v = i < n ? sums[i] : 0u;
        unsigned incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(SGB_FULL, incl, d); if (lane >= d) incl += u; }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            unsigned s = wsum[lane], si = s;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(SGB_FULL, si, d); if (lane >= d) si += u; }
            wsum[lane] = si - s;
        }
        __syncthreads();
        const unsigned ex = carry + wsum[wid] + incl - v;
        if (i < n) sums[i] = ex;
        __syncthreads();
        if (threadIdx.x == 1023) carry = ex + v;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total) *total = carry;
}
__global__ void __launch_bounds__(256) k_scan_add(unsigned* __restrict__ out, long long n, const unsigned* __restrict__ tile_sums)
{
    const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * 16;
    const unsigned off = tile_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < 16; i++) if (base + i < n) out[base + i] += off;
}
// out[i] = sum of in[0..i-1]; *d_total (device, optional) = sum of all.  tile_sums: ceil(n / 4096) words of scratch.
static cudaError_t scan_u32(const unsigned* in, unsigned* out, long long n, unsigned* tile_sums, unsigned long long* d_total, cudaStream_t st)
{
    if (n <= 0) { if (d_total) return cudaMemsetAsync(d_total, 0, 8, st); return cudaSuccess; }
    const int nt = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    k_scan_tiles<<<nt, 256, 0, st>>>(in, out, n, tile_sums);
    k_scan_sums<<<1, 1024, 0, st>>>(tile_sums, nt, d_total);
    k_scan_add<<<nt, 256, 0, st>>>(out, n, tile_sums);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ LSD radix sort of (uint64 key, uint32 value), stable
#define RS_TILE 2048                   // keys per block: 256 threads x 8
#define RS_BINS 256
__global__ void __launch_bounds__(256) k_rs_hist(const unsigned long long* __restrict__ keys, long long n, int shift,
                                                 unsigned* __restrict__ hist /* [RS_BINS][nblocks] */, int nblocks)
{
    __shared__ unsigned h[RS_BINS];
    h[threadIdx.x] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * RS_TILE;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const long long j = base + i * 256 + threadIdx.x;
        if (j < n) atomicAdd(&h[(unsigned)(keys[j] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(long long)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}
// One block scatters its tile in order: warp w owns elements [w * 256, (w + 1) * 256) of the tile, processed in 8 rounds of
// 32; the rank of an element among the tile's earlier elements with the same digit keeps the sort stable.  The tile is
// first put in digit order in shared memory and then written out by consecutive threads, so that the ~8 elements a tile holds
// of each digit leave as one contiguous run (a store straight from the registers touched up to 32 sectors per instruction with
// 8 or 4 useful bytes each).
__global__ void __launch_bounds__(256) k_rs_scatter(const unsigned long long* __restrict__ keys, const unsigned* __restrict__ vals,
                                                    long long n, int shift, const unsigned* __restrict__ offs /* hist scanned inside its scan tiles */,
                                                    const unsigned* __restrict__ tile_offs /* scanned sums of those tiles */,
                                                    int nblocks, unsigned long long* __restrict__ keys_out, unsigned* __restrict__ vals_out)
{
    __shared__ unsigned cnt[8][RS_BINS];          // per warp, per digit: elements of this warp, then their start inside the digit's run
    __shared__ unsigned lstart[RS_BINS];          // start of the digit's run inside the tile
    __shared__ unsigned gbase[RS_BINS];           // start of the digit's run in the output
    __shared__ unsigned wsum[8];
    __shared__ unsigned long long keyS[RS_TILE];
    __shared__ unsigned valS[RS_TILE];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8 * RS_BINS; i += 256) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const long long tile0 = (long long)blockIdx.x * RS_TILE;
    const long long tbase = tile0 + wid * 256;
    unsigned long long k[8];
    unsigned v[8];
    unsigned dg[8], rk[8];
    // pass 1: ranks inside the warp's 256 elements
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const long long j = tbase + r * 32 + lane;
        const bool ok = j < n;
        k[r] = ok ? keys[j] : 0ull;
        v[r] = ok ? vals[j] : 0u;
        dg[r] = ok ? ((unsigned)(k[r] >> shift) & 255u) : 0xffffffffu;
        const unsigned m = __match_any_sync(SGB_FULL, dg[r]);
        const unsigned before = __popc(m & ((1u << lane) - 1u));
        unsigned prior = 0;
        if (ok) prior = cnt[wid][dg[r]];
        __syncwarp();
        if (ok && before == 0) cnt[wid][dg[r]] = prior + __popc(m);      // the first lane of each digit class updates the count
        __syncwarp();
        rk[r] = prior + before;
    }
    __syncthreads();
    // digit d: its elements of warp w follow those of warps < w; its run starts after the runs of the smaller digits
    {
        const int d = threadIdx.x;
        unsigned tot = 0;
        for (int w = 0; w < 8; w++) { const unsigned c = cnt[w][d]; cnt[w][d] = tot; tot += c; }
        unsigned incl = tot;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const unsigned u = __shfl_up_sync(SGB_FULL, incl, s);
            if (lane >= s) incl += u;
        }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        unsigned woff = 0;
        for (int w = 0; w < wid; w++) woff += wsum[w];
        lstart[d] = woff + incl - tot;
        const long long hi = (long long)d * nblocks + blockIdx.x;
        gbase[d] = offs[hi] + tile_offs[hi / SCAN_TILE];        // (the scan's last pass, folded in: one launch less per digit)
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 8; r++) {
        if (dg[r] != 0xffffffffu) {
            const unsigned lp = lstart[dg[r]] + cnt[wid][dg[r]] + rk[r];
            keyS[lp] = k[r];
            valS[lp] = v[r];
        }
    }
    __syncthreads();
    const long long left = n - tile0;
    const int ntile = left < RS_TILE ? (int)left : RS_TILE;
    for (int i = threadIdx.x; i < ntile; i += 256) {
        const unsigned long long kk = keyS[i];
        const unsigned d = (unsigned)(kk >> shift) & 255u;
        const unsigned pos = gbase[d] + ((unsigned)i - lstart[d]);
        keys_out[pos] = kk;
        vals_out[pos] = valS[i];
    }
}
// sorts n (key, value) pairs by the key bits [bit_lo, bit_hi); ping-pongs between (k0, v0) and (k1, v1); returns in *which
// the index (0 / 1) of the buffers that hold the result.  hist: RS_BINS * nblocks words, tile_sums: scan scratch.
static cudaError_t radix_sort(unsigned long long* k0, unsigned* v0, unsigned long long* k1, unsigned* v1, long long n, int bit_lo,
                              int bit_hi, unsigned* hist, unsigned* tile_sums, int* which, cudaStream_t st)
{
    *which = 0;
    if (n <= 1) return cudaSuccess;
    const int nblocks = (int)((n + RS_TILE - 1) / RS_TILE);
    int cur = 0;
    for (int shift = bit_lo; shift < bit_hi; shift += 8) {
        unsigned long long* ki = cur ? k1 : k0; unsigned* vi = cur ? v1 : v0;
        unsigned long long* ko = cur ? k0 : k1; unsigned* vo = cur ? v0 : v1;
        k_rs_hist<<<nblocks, 256, 0, st>>>(ki, n, shift, hist, nblocks);
        const long long nh = (long long)RS_BINS * nblocks;
        const int nt = (int)((nh + SCAN_TILE - 1) / SCAN_TILE);
        k_scan_tiles<<<nt, 256, 0, st>>>(hist, hist, nh, tile_sums);
        k_scan_sums<<<1, 1024, 0, st>>>(tile_sums, nt, nullptr);
        k_rs_scatter<<<nblocks, 256, 0, st>>>(ki, vi, n, shift, hist, tile_sums, nblocks, ko, vo);
        cur ^= 1;
    }
    *which = cur;
    return cudaGetLastError();
}
static int bits_for(long long maxval) { int b = 1; while (b < 63 && (1ll << b) <= maxval) b++; return b; }

// ------------------------------------------------------------------ table order + combine
// key = cell << (cb + sb) | chr << sb | start  (cb = bits of the chromosome code, sb = bits of the largest start in the
// table: 28 for a human genome, which saves one 8-bit radix pass against a fixed 32); rows failing the filter (or with a bad
// key) get the DROP bit above the key bits and so sort last
__global__ void __launch_bounds__(256) k_tab_maxstart(const int32_t* __restrict__ start, long long n, int* __restrict__ out)
{
    int m = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int s = start[i];
        m = s > m ? s : m;
    }
    m = __reduce_max_sync(SGB_FULL, m);
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}
__global__ void __launch_bounds__(256) k_tab_keys(const int32_t* __restrict__ cell, const int32_t* __restrict__ chr,
                                                  const int32_t* __restrict__ start, const int32_t* __restrict__ total, double filtern,
                                                  int use_filter, long long n, int chr_bits, int start_bits, int drop_bit, int ncell, int nchr,
                                                  unsigned long long* __restrict__ keys, unsigned* __restrict__ idx, int* err)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c = cell[i], h = chr[i], s = start[i];
        unsigned long long key;
        const unsigned long long DROP = 1ull << drop_bit;
        if ((unsigned)c >= (unsigned)ncell || (unsigned)h >= (unsigned)nchr || s < 0) { atomicOr(err, SGB_ERR_BAD_KEY); key = DROP; }
        else key = ((unsigned long long)(unsigned)c << (chr_bits + start_bits)) | ((unsigned long long)(unsigned)h << start_bits) | (unsigned)s;
        if (use_filter && !((double)total[i] > filtern)) key |= DROP;             // .[totalcounts > filtern]
        keys[i] = key;
        idx[i] = (unsigned)i;
    }
}
// head[i] = 1 where a new (cell, chr, start) group starts among the sorted, kept rows
__global__ void __launch_bounds__(256) k_tab_heads(const unsigned long long* __restrict__ keys, long long n, int drop_bit,
                                                   unsigned* __restrict__ head)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long k = keys[i];
        head[i] = (!((k >> drop_bit) & 1ull) && (i == 0 || keys[i - 1] != k)) ? 1u : 0u;
    }
}
struct TabCols { const int32_t *cell, *chr, *start, *state, *a, *b, *n; };
// the four payload columns of a row as one 16-byte record: the gather through the sort permutation then costs one sector per
// row instead of four (each of which moved 32 bytes for 4 useful ones; with cell / chr / start gathered too the first version
// read 18.7 GB for 22.8 M rows)
__global__ void __launch_bounds__(256) k_tab_pack(TabCols C, long long n, int4* __restrict__ recs)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        recs[i] = make_int4(C.state[i], C.a[i], C.b[i], C.n[i]);
}
// one thread per group head: sums its run (stable sort: the run is in input order), writes the group's row, counts bins.
// cell / chr / start come out of the sorted key; the per-cell and per-(cell, chr) counters are bumped once per warp and key
// (the rows are sorted by cell and chromosome, so a warp's heads mostly share them).
__global__ void __launch_bounds__(256) k_tab_reduce(const unsigned long long* __restrict__ keys, const unsigned* __restrict__ idx,
                                                    const unsigned* __restrict__ head, const unsigned* __restrict__ gpos, long long n,
                                                    const int4* __restrict__ recs, int nchr, int chr_bits, int start_bits,
                                                    int32_t* g_first, int32_t* g_cell, int32_t* g_chr,
                                                    int32_t* g_start, int32_t* g_state, int32_t* g_a, int32_t* g_b, int32_t* g_n,
                                                    unsigned* cnt_cell, unsigned* cnt_cc, int* err)
{
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nround = (n + stride - 1) / stride;
    for (long long it = 0; it < nround; it++) {
        const long long i = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const bool is_head = i < n && head[i];
        unsigned long long cc = ~0ull;             // (cell, chr) of this lane's group, ~0 when the lane has none
        if (is_head) {
            const unsigned long long k = keys[i];
            const unsigned r0 = idx[i];
            const int4 f = recs[r0];
            long long sa = 0, sb = 0, sn = 0;
            for (long long j = i; j < n && keys[j] == k; j++) {
                const int4 q = j == i ? f : recs[idx[j]];
                sa += q.y; sb += q.z; sn += q.w;
                if (q.x != f.x) atomicOr(err, SGB_ERR_KEY_CONFLICT);       // (chr, start, cell) must determine state (and end, copy)
            }
            if (sa > 0x7fffffffll || sb > 0x7fffffffll || sn > 0x7fffffffll || sa < 0 || sb < 0 || sn < 0) atomicOr(err, SGB_ERR_COUNT_TOO_LARGE);
            const unsigned g = gpos[i];
            const int c = (int)(k >> (chr_bits + start_bits)), h = (int)((k >> start_bits) & ((1ull << chr_bits) - 1ull));
            g_first[g] = (int32_t)r0; g_cell[g] = c; g_chr[g] = h; g_start[g] = (int32_t)(k & ((1ull << start_bits) - 1ull));
            g_state[g] = f.x; g_a[g] = (int32_t)sa; g_b[g] = (int32_t)sb; g_n[g] = (int32_t)sn;
            cc = k >> start_bits;
        }
        const unsigned peers = __match_any_sync(SGB_FULL, cc);
        if (is_head && lane == __ffs(peers) - 1) {
            const int c = (int)(cc >> chr_bits), h = (int)(cc & ((1ull << chr_bits) - 1ull));
            atomicAdd(&cnt_cell[c], (unsigned)__popc(peers));
            atomicAdd(&cnt_cc[(long long)c * nchr + h], (unsigned)__popc(peers));
        }
    }
}
// per cell: n > minbins and min over its chromosomes of (chr == X ? 100 : n_chr) > minbinschr   (R/combinedata.R:314-323)
__global__ void __launch_bounds__(256) k_tab_cellkeep(const unsigned* __restrict__ cnt_cell, const unsigned* __restrict__ cnt_cc, int ncell,
                                                      int nchr, int chrX, int minbins, int minbinschr, unsigned char* __restrict__ keep)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    const long long nc = cnt_cell[c];
    long long mn = 0x7fffffff;
    bool any = false;
    for (int h = 0; h < nchr; h++) {
        const long long v = cnt_cc[(long long)c * nchr + h];
        if (v == 0) continue;
        const long long w = (h == chrX) ? 100 : v;
        mn = w < mn ? w : mn;
        any = true;
    }
    keep[c] = (unsigned char)(any && nc > minbins && mn > minbinschr);
}
__global__ void __launch_bounds__(256) k_tab_keepflags(const int32_t* __restrict__ g_cell, const int32_t* __restrict__ g_chr, long long ng,
                                                       const unsigned char* __restrict__ keep, unsigned* __restrict__ kflag,
                                                       unsigned* __restrict__ sflag)
{
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += (long long)gridDim.x * blockDim.x) {
        const int c = g_cell[g];
        const unsigned k = keep[c];
        kflag[g] = k;
        // a kept group starts a (cell, chr) sequence when its predecessor differs in cell or chromosome (cells are kept or
        // dropped as a whole, so the predecessor of a kept group inside the same cell is kept too)
        sflag[g] = (k && (g == 0 || g_cell[g - 1] != c || g_chr[g - 1] != g_chr[g])) ? 1u : 0u;
    }
}
struct TabOut { int32_t *first, *cell, *chr, *start, *state, *a, *b, *n; };
__global__ void __launch_bounds__(256) k_tab_compact(long long ng, const unsigned* __restrict__ kflag, const unsigned* __restrict__ kpos,
                                                     const unsigned* __restrict__ sflag, const unsigned* __restrict__ spos, TabOut I, TabOut O,
                                                     long long* __restrict__ seg_off)
{
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += (long long)gridDim.x * blockDim.x) {
        if (!kflag[g]) continue;
        const unsigned o = kpos[g];
        O.first[o] = I.first[g]; O.cell[o] = I.cell[g]; O.chr[o] = I.chr[g]; O.start[o] = I.start[g]; O.state[o] = I.state[g];
        O.a[o] = I.a[g]; O.b[o] = I.b[g]; O.n[o] = I.n[g];
        if (sflag[g]) seg_off[spos[g]] = (long long)o;
    }
}
// table order only: perm = idx, sequence heads where (cell, chr) changes
__global__ void __launch_bounds__(256) k_tab_seqheads(const unsigned long long* __restrict__ keys, long long n, int start_bits,
                                                      unsigned* __restrict__ sflag)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        sflag[i] = (i == 0 || (keys[i - 1] >> start_bits) != (keys[i] >> start_bits)) ? 1u : 0u;
}
__global__ void __launch_bounds__(256) k_tab_segwrite(const unsigned* __restrict__ sflag, const unsigned* __restrict__ spos, long long n,
                                                      long long* __restrict__ seg_off)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        if (sflag[i]) seg_off[spos[i]] = i;
}

static unsigned grid_for(long long n, int per = 256, int cap = 148 * 16)
{
    long long b = (n + per - 1) / per;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

size_t sgbk_table_scratch_bytes(long long n, int ncell, int nchr)
{
    const size_t N = (size_t)(n > 0 ? n : 1);
    const size_t nb = (N + RS_TILE - 1) / RS_TILE;
    size_t b = 0;
    b += 2 * N * 8 + 2 * N * 4;                                   // keys, idx (ping-pong)
    b += 4 * N * 4;                                               // head / flags / positions (4 uint32 arrays)
    b += 8 * N * 4;                                               // group rows (8 int32 columns)
    b += N * 16 + 256;                                            // packed payload records
    b += (size_t)RS_BINS * nb * 4 + ((size_t)RS_BINS * nb / SCAN_TILE + 2 + N / SCAN_TILE + 2) * 4;
    b += ((size_t)ncell + (size_t)ncell * nchr) * 4 + (size_t)ncell;
    return b + 32 * 256 + 4096;
}

struct Carver {
    char* p; size_t off;
    template <typename T> T* take(size_t count) { off = (off + 255) & ~(size_t)255; T* r = (T*)(p + off); off += count * sizeof(T); return r; }
};

// The sort of the long table by (cell, chr, start) and, with do_combine, combineBAFCN's keyed reduce and filters.
// d_* inputs: device columns of n rows.  Outputs: device arrays of capacity n (seg_off: n + 1).  Synchronises the stream
// twice (the number of groups and of kept rows are only known on the device).  h_counts[0] = rows out, [1] = sequences.
// do_combine = 0: o_first receives the sort permutation (no reduce, no filters); the other outputs are not written.
cudaError_t sgbk_table(long long n, const int32_t* d_cell, const int32_t* d_chr, const int32_t* d_start, const int32_t* d_state,
                       const int32_t* d_a, const int32_t* d_b, const int32_t* d_n, int ncell, int nchr, int do_combine, double filtern,
                       int minbins, int minbinschr, int chrX, void* d_scratch, int32_t* o_first, int32_t* o_cell, int32_t* o_chr,
                       int32_t* o_start, int32_t* o_state, int32_t* o_a, int32_t* o_b, int32_t* o_n, long long* o_seg_off,
                       long long* h_counts, int* d_err, cudaStream_t st)
{
    h_counts[0] = 0; h_counts[1] = 0;
    if (n <= 0) return cudaSuccess;
    const size_t N = (size_t)n;
    const size_t nb = (N + RS_TILE - 1) / RS_TILE;
    Carver cv{(char*)d_scratch, 0};
    unsigned long long* k0 = cv.take<unsigned long long>(N); unsigned long long* k1 = cv.take<unsigned long long>(N);
    unsigned* v0 = cv.take<unsigned>(N); unsigned* v1 = cv.take<unsigned>(N);
    unsigned* f0 = cv.take<unsigned>(N); unsigned* f1 = cv.take<unsigned>(N); unsigned* f2 = cv.take<unsigned>(N); unsigned* f3 = cv.take<unsigned>(N);
    TabOut G;
    G.first = cv.take<int32_t>(N); G.cell = cv.take<int32_t>(N); G.chr = cv.take<int32_t>(N); G.start = cv.take<int32_t>(N);
    G.state = cv.take<int32_t>(N); G.a = cv.take<int32_t>(N); G.b = cv.take<int32_t>(N); G.n = cv.take<int32_t>(N);
    unsigned* hist = cv.take<unsigned>((size_t)RS_BINS * nb);
    unsigned* tsum = cv.take<unsigned>((size_t)RS_BINS * nb / SCAN_TILE + 2 + N / SCAN_TILE + 2);
    unsigned* cnt_cell = cv.take<unsigned>((size_t)ncell + (size_t)ncell * nchr);
    unsigned* cnt_cc = cnt_cell + ncell;
    unsigned char* keep = cv.take<unsigned char>((size_t)ncell);
    unsigned long long* d_tot = cv.take<unsigned long long>(4);
    int4* recs = do_combine ? cv.take<int4>(N) : nullptr;
    const int chr_bits = bits_for(nchr - 1 > 0 ? nchr - 1 : 1);
    cudaError_t e;
    // the largest start decides how many key bits (and radix passes) the position needs
    int* d_maxstart = (int*)(d_tot + 3);
    int maxstart = 0;
    if ((e = cudaMemsetAsync(d_maxstart, 0, sizeof(int), st)) != cudaSuccess) return e;
    k_tab_maxstart<<<grid_for(n), 256, 0, st>>>(d_start, n, d_maxstart);
    if ((e = cudaMemcpyAsync(&maxstart, d_maxstart, sizeof(int), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    const int start_bits = bits_for(maxstart > 1 ? maxstart : 1);
    const int key_bits = start_bits + chr_bits + bits_for(ncell - 1 > 0 ? ncell - 1 : 1);
    if (key_bits > 62) return cudaErrorInvalidValue;
    const int drop_bit = key_bits;
    k_tab_keys<<<grid_for(n), 256, 0, st>>>(d_cell, d_chr, d_start, d_n, filtern, do_combine, n, chr_bits, start_bits, drop_bit, ncell, nchr, k0, v0, d_err);
    int which = 0;
    if ((e = radix_sort(k0, v0, k1, v1, n, 0, key_bits + 1, hist, tsum, &which, st)) != cudaSuccess) return e;
    unsigned long long* ks = which ? k1 : k0; unsigned* vs = which ? v1 : v0;
    if (!do_combine) {
        if ((e = cudaMemcpyAsync(o_first, vs, N * 4, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        k_tab_seqheads<<<grid_for(n), 256, 0, st>>>(ks, n, start_bits, f0);
        if ((e = scan_u32(f0, f1, n, tsum, d_tot, st)) != cudaSuccess) return e;
        k_tab_segwrite<<<grid_for(n), 256, 0, st>>>(f0, f1, n, o_seg_off);
        unsigned long long nseg = 0;
        if ((e = cudaMemcpyAsync(&nseg, d_tot, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
        const long long end = n;
        if ((e = cudaMemcpyAsync(o_seg_off + nseg, &end, 8, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
        h_counts[0] = n; h_counts[1] = (long long)nseg;
        return cudaGetLastError();
    }
    k_tab_heads<<<grid_for(n), 256, 0, st>>>(ks, n, drop_bit, f0);
    if ((e = scan_u32(f0, f1, n, tsum, d_tot, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(cnt_cell, 0, ((size_t)ncell + (size_t)ncell * nchr) * 4, st)) != cudaSuccess) return e;
    TabCols C{d_cell, d_chr, d_start, d_state, d_a, d_b, d_n};
    k_tab_pack<<<grid_for(n), 256, 0, st>>>(C, n, recs);
    k_tab_reduce<<<grid_for(n), 256, 0, st>>>(ks, vs, f0, f1, n, recs, nchr, chr_bits, start_bits, G.first, G.cell, G.chr, G.start, G.state,
                                              G.a, G.b, G.n, cnt_cell, cnt_cc, d_err);
    k_tab_cellkeep<<<(ncell + 255) / 256, 256, 0, st>>>(cnt_cell, cnt_cc, ncell, nchr, chrX, minbins, minbinschr, keep);
    unsigned long long ng = 0;
    if ((e = cudaMemcpyAsync(&ng, d_tot, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    if (ng == 0) return cudaGetLastError();
    k_tab_keepflags<<<grid_for((long long)ng), 256, 0, st>>>(G.cell, G.chr, (long long)ng, keep, f2, f3);
    if ((e = scan_u32(f2, f0, (long long)ng, tsum, d_tot, st)) != cudaSuccess) return e;
    if ((e = scan_u32(f3, f1, (long long)ng, tsum, d_tot + 1, st)) != cudaSuccess) return e;
    TabOut O{o_first, o_cell, o_chr, o_start, o_state, o_a, o_b, o_n};
    k_tab_compact<<<grid_for((long long)ng), 256, 0, st>>>((long long)ng, f2, f0, f3, f1, G, O, o_seg_off);
    unsigned long long tot[2] = {0, 0};
    if ((e = cudaMemcpyAsync(tot, d_tot, 16, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    const long long end = (long long)tot[0];
    if ((e = cudaMemcpyAsync(o_seg_off + tot[1], &end, 8, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    h_counts[0] = (long long)tot[0]; h_counts[1] = (long long)tot[1];
    return cudaGetLastError();
}

// ------------------------------------------------------------------ between-pass sums
// Rows are the decoded table; A, B from the clamp block (R/callHSCN.R:282-290); cell_of_seq / keep select the cells that
// survived the first-pass filter (R/callHSCN.R:883-893: ascn_filt).  Exact integer sums (128-bit where squares are summed):
//   out[0..2]  sum alleleB, sum totalcounts, rows  over A-Hom rows (B == 0 < A)          -> infloherror (:895-899)
//   out[3..8]  over the rows in state (selA | selB): rows, sum m, sum n, sum m^2, sum m n, sum n^2 with m = alleleB,
//              n = totalcounts                                                          -> tarones_Z (:609-616)
//   hist[A * S + B]  rows with B > 0                                                    -> fitBB's modal state (:619)
__global__ void __launch_bounds__(256) k_pass_stats2(const int32_t* __restrict__ n, const int32_t* __restrict__ x, const void* __restrict__ cn,
                                                     int cn_itemsize, const uint8_t* __restrict__ path, const uint8_t* __restrict__ rowkeep,
                                                     long long N, int S, int selA, int selB, double* __restrict__ out9,
                                                     unsigned long long* __restrict__ hist_out)
{
    __shared__ unsigned int hist[SGB_MAX_STATES * SGB_MAX_STATES];
    __shared__ double red[8][9];
    for (int i = threadIdx.x; i < S * S; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    unsigned long long hx = 0, hn = 0, hc = 0, tc = 0, tm = 0, tn = 0;
    unsigned __int128 tmm = 0, tmn = 0, tnn = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        if (rowkeep && !rowkeep[i]) continue;
        const int st = (cn_itemsize == 1) ? (int)((const uint8_t*)cn)[i] : ((const int32_t*)cn)[i];
        const int sm = path[i];
        int a = st - sm, b = sm;
        if (a < 0) a = 0;
        if (b > st) b = st;
        if (a > st) a = st;
        const unsigned long long m = (unsigned long long)x[i], nn = (unsigned long long)n[i];
        if (b == 0 && a > 0) { hx += m; hn += nn; hc++; }
        if (b > 0 && a < S && b < S) atomicAdd(&hist[a * S + b], 1u);
        if (a == selA && b == selB) { tc++; tm += m; tn += nn; tmm += (unsigned __int128)m * m; tmn += (unsigned __int128)m * nn; tnn += (unsigned __int128)nn * nn; }
    }
    // per-thread exact integers -> doubles (each conversion correctly rounded), then a fixed-order tree: deterministic
    double v[9] = {(double)hx, (double)hn, (double)hc, (double)tc, (double)tm, (double)tn, (double)tmm, (double)tmn, (double)tnn};
#pragma unroll
    for (int k = 0; k < 9; k++)
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(SGB_FULL, v[k], off);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) for (int k = 0; k < 9; k++) red[wid][k] = v[k];
    __syncthreads();
    if (threadIdx.x < 9) {
        double s = 0.0;
        for (int w = 0; w < 8; w++) s += red[w][threadIdx.x];
        out9[(long long)blockIdx.x * 9 + threadIdx.x] = s;
    }
    for (int i = threadIdx.x; i < S * S; i += blockDim.x)
        if (hist[i]) atomicAdd(&hist_out[i], (unsigned long long)hist[i]);
}
__global__ void k_pass_stats2_final(const double* __restrict__ part, int nblocks, double* __restrict__ out9)
{
    if (threadIdx.x < 9) {
        double s = 0.0;
        for (int b = 0; b < nblocks; b++) s += part[(long long)b * 9 + threadIdx.x];
        out9[threadIdx.x] = s;
    }
}
int sgbk_pass_stats2_blocks(long long N) { return (int)grid_for(N, 256, 148 * 4); }
cudaError_t sgbk_pass_stats2(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize, const uint8_t* d_path,
                             const uint8_t* d_rowkeep, long long N, int S, int selA, int selB, double* d_part, double* d_out9,
                             unsigned long long* d_hist, cudaStream_t st)
{
    cudaError_t e = cudaMemsetAsync(d_hist, 0, (size_t)S * S * 8, st);
    if (e != cudaSuccess) return e;
    const int nb = sgbk_pass_stats2_blocks(N);
    k_pass_stats2<<<nb, 256, 0, st>>>(d_n, d_x, d_cn, cn_itemsize, d_path, d_rowkeep, N, S, selA, selB, d_part, d_hist);
    k_pass_stats2_final<<<1, 32, 0, st>>>(d_part, nb, d_out9);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ qc_per_cell (R/util.R:667-691)
// key per row: (cell, A, B) with rows in state 0|0 dropped; value: BAF = alleleB / totalcounts as its IEEE bit pattern
// (non-negative doubles order like their bits; NaN rows -- totalcounts 0 -- are dropped: median(..., na.rm = TRUE)).
// Two stable radix sorts (by BAF bits, then by key) leave every (cell, A, B) group contiguous with its BAFs ascending;
// the group head takes the median, and per cell the distances |B/(A+B) - median| of the groups with n > 10 are averaged
// with weights n / Ntot (Ntot = rows outside 0|0, all cells) exactly as weighted.mean(distance, frac) does.
__global__ void __launch_bounds__(256) k_qc_keys(const int32_t* __restrict__ n, const int32_t* __restrict__ x, const void* __restrict__ cn,
                                                 int cn_itemsize, const uint8_t* __restrict__ path, const int32_t* __restrict__ cell_of_row,
                                                 long long N, int ncell, int drop_bit, unsigned long long* __restrict__ kbaf,
                                                 unsigned long long* __restrict__ kgrp, unsigned* __restrict__ idx,
                                                 unsigned long long* __restrict__ ntot, int* err)
{
    unsigned long long mine = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const int st = (cn_itemsize == 1) ? (int)((const uint8_t*)cn)[i] : ((const int32_t*)cn)[i];
        const int sm = path[i];
        int a = st - sm, b = sm;
        if (a < 0) a = 0;
        if (b > st) b = st;
        if (a > st) a = st;
        const int c = cell_of_row[i];
        if ((unsigned)c >= (unsigned)ncell || (unsigned)st > 255u) atomicOr(err, SGB_ERR_BAD_KEY);
        const double baf = sgb_div((double)x[i], (double)n[i]);
        unsigned long long g = ((unsigned long long)(unsigned)c << 16) | (unsigned long long)((a & 255) * 256 + (b & 255));
        const bool zero = (a == 0 && b == 0);                          // state_AS_phased == "0|0": filtered out
        if (!zero) mine++;
        if (zero || baf != baf) g |= 1ull << drop_bit;                 // (BAF is NA only when totalcounts == 0)
        kbaf[i] = baf != baf ? 0ull : sgb_d2u(baf);
        kgrp[i] = g;
        idx[i] = (unsigned)i;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(SGB_FULL, mine, off);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(ntot, mine);
}
__global__ void __launch_bounds__(256) k_qc_gather(const unsigned long long* __restrict__ kgrp, const unsigned* __restrict__ idx, long long N,
                                                   unsigned long long* __restrict__ out)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) out[i] = kgrp[idx[i]];
}
// kg: group keys sorted (rows of a cell contiguous, groups inside it contiguous, BAF ascending inside a group);
// rows[i]: the table row at sorted position i.  One WARP per cell: its row range comes from two binary searches on the
// sorted keys, the lanes scan it 32 rows at a time for group boundaries, and the (few) groups are closed in order with the
// arithmetic of the reference -- every lane computes the same values.  (One thread per cell walking ~5,700 rows took 2.5 ms
// for 4,000 cells.)
__device__ __forceinline__ long long qc_lower_bound(const unsigned long long* __restrict__ kg, long long N, unsigned long long cellkey)
{
    long long lo = 0, hi = N;                       // first i with (kg[i] >> 16) >= cellkey; dropped rows carry a bit above the cell bits
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if ((kg[mid] >> 16) < cellkey) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__global__ void __launch_bounds__(128) k_qc_cells(const unsigned long long* __restrict__ kg, const unsigned* __restrict__ rows,
                                                  const unsigned long long* __restrict__ kbaf, long long N, int ncell,
                                                  const unsigned long long* __restrict__ ntot, double* __restrict__ avg)
{
    const double ntd = (double)*ntot;
    const int lane = threadIdx.x & 31;
    const int cell = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (cell >= ncell) return;
    long long s = 0;
    if (lane < 2) s = qc_lower_bound(kg, N, (unsigned long long)cell + lane);
    const long long e = __shfl_sync(SGB_FULL, s, 1);
    s = __shfl_sync(SGB_FULL, s, 0);
    if (s >= e) return;                                                  // no rows: avg[cell] stays NaN
    double num = 0.0, den = 0.0;
    long long gstart = s;                                                // first row of the open group
    unsigned long long gkey = kg[s];
    auto close_group = [&](long long gend) {
        const long long cnt = gend - gstart;
        if (cnt > 10) {                                                  // dplyr::filter(n > 10)
            const double lo = sgb_u2d(kbaf[rows[gstart + (cnt - 1) / 2]]), hi = sgb_u2d(kbaf[rows[gstart + cnt / 2]]);
            const double med = sgb_mul(sgb_add(lo, hi), 0.5);            // R's median: mean of the two middle values
            const int ab = (int)(gkey & 0xffffull), a = ab >> 8, b = ab & 255;
            const double d = sgb_fabs(sgb_sub(sgb_div((double)b, (double)(a + b)), med));      // sqrt((expBAF - medianBAF)^2)
            const double frac = sgb_div((double)cnt, ntd);
            num = sgb_add(num, sgb_mul(d, frac));
            den = sgb_add(den, frac);
        }
    };
    for (long long j0 = s + 1; j0 < e; j0 += 32) {
        const long long j = j0 + lane;
        const bool headj = j < e && kg[j] != kg[j - 1];
        unsigned m = __ballot_sync(SGB_FULL, headj);
        while (m) {                                                       // group heads of this stretch, in row order
            const int l = __ffs(m) - 1;
            m &= m - 1;
            const long long p = j0 + l;
            close_group(p);
            gstart = p;
            gkey = kg[p];
        }
    }
    close_group(e);
    if (lane == 0) avg[cell] = sgb_div(num, den);                         // weighted.mean(distance, frac); NaN when no state qualifies
}
size_t sgbk_qc_scratch_bytes(long long n)
{
    const size_t N = (size_t)(n > 0 ? n : 1);
    const size_t nb = (N + RS_TILE - 1) / RS_TILE;
    return 4 * N * 8 + 2 * N * 4 + (size_t)RS_BINS * nb * 4 + ((size_t)RS_BINS * nb / SCAN_TILE + 4) * 4 + 16 * 256 + 4096;
}
// d_avg (ncell doubles) is filled with NaN first: cells without any state of more than 10 bins stay NaN
cudaError_t sgbk_qc_per_cell(const int32_t* d_n, const int32_t* d_x, const void* d_cn, int cn_itemsize, const uint8_t* d_path,
                             const int32_t* d_cell_of_row, long long N, int ncell, void* d_scratch, double* d_avg, int* d_err, cudaStream_t st)
{
    cudaError_t e;
    if ((e = cudaMemsetAsync(d_avg, 0xff, (size_t)ncell * 8, st)) != cudaSuccess) return e;      // all-ones = a NaN
    if (N <= 0) return cudaSuccess;
    const int cell_bits = bits_for(ncell - 1 > 0 ? ncell - 1 : 1);
    if (cell_bits + 16 > 62) return cudaErrorInvalidValue;
    const int drop_bit = 16 + cell_bits;
    const size_t Nn = (size_t)N, nb = (Nn + RS_TILE - 1) / RS_TILE;
    Carver cv{(char*)d_scratch, 0};
    unsigned long long* kbaf = cv.take<unsigned long long>(Nn); unsigned long long* kgrp = cv.take<unsigned long long>(Nn);
    unsigned long long* k0 = cv.take<unsigned long long>(Nn); unsigned long long* k1 = cv.take<unsigned long long>(Nn);
    unsigned* v0 = cv.take<unsigned>(Nn); unsigned* v1 = cv.take<unsigned>(Nn);
    unsigned* hist = cv.take<unsigned>((size_t)RS_BINS * nb);
    unsigned* tsum = cv.take<unsigned>((size_t)RS_BINS * nb / SCAN_TILE + 4);
    unsigned long long* ntot = cv.take<unsigned long long>(2);
    if ((e = cudaMemsetAsync(ntot, 0, 16, st)) != cudaSuccess) return e;
    k_qc_keys<<<grid_for(N), 256, 0, st>>>(d_n, d_x, d_cn, cn_itemsize, d_path, d_cell_of_row, N, ncell, drop_bit, kbaf, kgrp, v0, ntot, d_err);
    // sort 1: rows by BAF (a copy of the BAF bits is the key; kbaf itself stays in row order for the look-ups)
    if ((e = cudaMemcpyAsync(k0, kbaf, Nn * 8, cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
    int w = 0;
    if ((e = radix_sort(k0, v0, k1, v1, N, 0, 64, hist, tsum, &w, st)) != cudaSuccess) return e;
    unsigned* ord1 = w ? v1 : v0; unsigned* ordt = w ? v0 : v1;
    unsigned long long* ka = w ? k0 : k1;           // free buffer for the gathered group keys
    unsigned long long* kb = w ? k1 : k0;
    k_qc_gather<<<grid_for(N), 256, 0, st>>>(kgrp, ord1, N, ka);
    // sort 2 (stable): by group key, values = table rows in BAF order
    int w2 = 0;
    if ((e = radix_sort(ka, ord1, kb, ordt, N, 0, drop_bit + 1, hist, tsum, &w2, st)) != cudaSuccess) return e;
    const unsigned long long* kg = w2 ? kb : ka; const unsigned* rows = w2 ? ordt : ord1;
    k_qc_cells<<<(unsigned)((ncell + 3) / 4), 128, 0, st>>>(kg, rows, kbaf, N, ncell, ntot, d_avg);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ fix_assignments (R/callHSCN.R:1072-1286), as written
// density (NOT its logarithm, as the reference calls dbinom / dbetabinom without log = TRUE) of k successes out of k + nk
__device__ double fix_density(double kd, double nkd, double p, int lik_bb, double rho, double bb_s)
{
    if (p != p || p < 0.0 || p > 1.0) return sgb_nan();
    const double n = sgb_add(kd, nkd);
    if (p == 0.0) return kd == 0.0 ? 1.0 : 0.0;
    if (p == 1.0) return nkd == 0.0 ? 1.0 : 0.0;
    const double lc = sgb_sub(sgb_sub(sgb_logfact((int64_t)n, gt_lftab, gt_logtab), sgb_logfact((int64_t)kd, gt_lftab, gt_logtab)),
                              sgb_logfact((int64_t)nkd, gt_lftab, gt_logtab));
    double l;
    const double big = 485165195.40979028;
    const double a = sgb_mul(p, bb_s), b = sgb_mul(sgb_sub(1.0, p), bb_s);
    if (!lik_bb || !(rho > 0.0) || (a > big && b > big)) {
        l = sgb_fma(nkd, sgb_log1p(-p, gt_logtab), sgb_fma(kd, sgb_log(p, gt_logtab), lc));
    } else {
        double t1 = sgb_add(sgb_lgamma(sgb_add(a, kd), gt_logtab), sgb_lgamma(sgb_add(b, nkd), gt_logtab));
        t1 = sgb_sub(t1, sgb_lgamma(sgb_add(bb_s, n), gt_logtab));
        const double t0 = sgb_sub(sgb_add(sgb_lgamma(a, gt_logtab), sgb_lgamma(b, gt_logtab)), sgb_lgamma(sgb_add(a, b), gt_logtab));
        l = sgb_add(lc, sgb_sub(t1, t0));
    }
    return sgb_exp(l, gt_exptab);
}
__device__ __forceinline__ double fix_p(int num, int state, double loh)
{
    double p = sgb_div((double)num, (double)state);
    if (p != p) p = 0.0;                                   // ifelse(is.nan(B / state), 0.0, B / state)
    if (p == 0.0) p = sgb_add(p, loh);
    if (p == 1.0) p = sgb_sub(p, loh);
    return p;
}
__device__ __forceinline__ void fix_clamp(int state, int smin, int& A, int& B)
{
    A = state - smin; B = smin;
    if (B < 0) B = 0;
    if (A < 0) A = 0;
    if (B > state) B = state;
    if (A > state) A = state;
}
// pass 1: singletons.  rleid(state_AS_phased) by (cell, chr): a row is a run of length 1 when its (A, B) differs from the
// neighbours inside its sequence.  shift(A) / shift(B, type = "lead") are taken over the WHOLE table (no `by`), fill = 0.
__global__ void __launch_bounds__(256) k_fix_pass1(const int32_t* __restrict__ state, const int32_t* __restrict__ A, const int32_t* __restrict__ B,
                                                   const int32_t* __restrict__ alleleA, const int32_t* __restrict__ alleleB,
                                                   const uint8_t* __restrict__ seqhead, long long N, double loh, int lik_bb, double rho,
                                                   double bb_s, int32_t* __restrict__ A1, int32_t* __restrict__ B1)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const int a = A[i], b = B[i], st = state[i];
        const bool first = seqhead[i] != 0, last = (i + 1 == N) || seqhead[i + 1] != 0;
        const bool same_prev = !first && A[i - 1] == a && B[i - 1] == b;
        const bool same_next = !last && A[i + 1] == a && B[i + 1] == b;
        const bool single = !same_prev && !same_next;                   // nbins == 1
        const int Bprev = i > 0 ? B[i - 1] : 0, Bafter = i + 1 < N ? B[i + 1] : 0;
        int smin = b;
        if (single) {
            const double kd = (double)alleleA[i], nkd = (double)alleleB[i];
            double LLafter = fix_density(kd, nkd, fix_p(Bafter, st, loh), lik_bb, rho, bb_s);
            if (LLafter != LLafter) LLafter = 0.0;                      // fifelse(is.na(LLafter), 0, LLafter)
            double LLprev = fix_density(kd, nkd, fix_p(Bprev, st, loh), lik_bb, rho, bb_s);
            LLprev = (LLprev != LLprev) ? 0.0 : LLafter;                // fifelse(is.na(LLprev), 0, LLafter)   [sic]
            // state_min := fifelse(nbins == 1 & LLafter < LLprev, Bprev, B) is overwritten by the next assignment:
            smin = (LLafter >= LLprev) ? Bafter : b;
        }
        int na, nb2;
        fix_clamp(st, smin, na, nb2);
        A1[i] = na; B1[i] = nb2;
    }
}
// pass 2: runs of equal (A, B) -- inside a sequence (binomial branch) or across the whole table (beta-binomial branch,
// rleid without `by`) -- totals of alleleA / alleleB over the run, swap A and B where the other assignment is likelier.
// Row-parallel: run heads are flagged and numbered by a scan, every row adds its counts to its run's totals (exact integer
// sums; one atomic per run and warp after a segmented shuffle reduction), then every row evaluates its own two densities.
// (The first version let the thread of a run head walk its run -- a run is most of a chromosome, so one lane in 32 worked
// through ~450 rows of lgamma-class arithmetic alone: 84 ms for 22.8 M rows.)
__global__ void __launch_bounds__(256) k_fix_runheads(const int32_t* __restrict__ A1, const int32_t* __restrict__ B1,
                                                      const uint8_t* __restrict__ seqhead, long long N, int global_runs,
                                                      unsigned* __restrict__ flag)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const bool brk = !global_runs && seqhead[i] != 0;
        flag[i] = (i == 0 || brk || A1[i - 1] != A1[i] || B1[i - 1] != B1[i]) ? 1u : 0u;
    }
}
// rid[i] = index of row i's run (flag scanned exclusively into pos); the head of a run clears its totals
__global__ void __launch_bounds__(256) k_fix_runids(const unsigned* __restrict__ flag, unsigned* __restrict__ pos_then_rid, long long N,
                                                    unsigned long long* __restrict__ tot)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const unsigned r = pos_then_rid[i] + flag[i] - 1u;
        pos_then_rid[i] = r;
        if (flag[i]) { tot[2ull * r] = 0ull; tot[2ull * r + 1] = 0ull; }
    }
}
__global__ void __launch_bounds__(256) k_fix_runtotals(const int32_t* __restrict__ alleleA, const int32_t* __restrict__ alleleB,
                                                       const unsigned* __restrict__ rid, long long N, unsigned long long* __restrict__ tot)
{
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nround = (N + stride - 1) / stride;
    for (long long it = 0; it < nround; it++) {
        const long long i = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const bool in = i < N;
        const unsigned r = in ? rid[i] : 0xffffffffu;
        long long va = in ? (long long)alleleA[i] : 0, vb = in ? (long long)alleleB[i] : 0;
        // rows of a run are consecutive, so the lanes of a run are: segmented reduction towards the run's first lane
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long ua = __shfl_down_sync(SGB_FULL, va, d), ub = __shfl_down_sync(SGB_FULL, vb, d);
            const unsigned ur = __shfl_down_sync(SGB_FULL, r, d);
            if (lane + d < 32 && ur == r) { va += ua; vb += ub; }
        }
        const unsigned rprev = __shfl_up_sync(SGB_FULL, r, 1);
        if (in && (lane == 0 || rprev != r)) {
            atomicAdd(&tot[2ull * r], (unsigned long long)va);
            atomicAdd(&tot[2ull * r + 1], (unsigned long long)vb);
        }
    }
}
__global__ void __launch_bounds__(256) k_fix_pass2(const int32_t* __restrict__ state, const int32_t* __restrict__ A1, const int32_t* __restrict__ B1,
                                                   const unsigned* __restrict__ rid, const unsigned long long* __restrict__ tot, long long N,
                                                   double loh, int lik_bb, double rho, double bb_s, int32_t* __restrict__ A2, int32_t* __restrict__ B2)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const int a = A1[i], b = B1[i], st = state[i];
        const unsigned r = rid[i];
        const double kd = (double)(long long)tot[2ull * r], nkd = (double)(long long)tot[2ull * r + 1];
        const double LLassigned = fix_density(kd, nkd, fix_p(b, st, loh), lik_bb, rho, bb_s);
        const double LLother = fix_density(kd, nkd, fix_p(a, st, loh), lik_bb, rho, bb_s);
        const int smin = (LLother < LLassigned) ? a : b;             // fifelse(LLother < LLassigned, A, B)
        int na, nb2;
        fix_clamp(st, smin, na, nb2);
        A2[i] = na; B2[i] = nb2;
    }
}
size_t sgbk_fix_scratch_bytes(long long N)
{
    const size_t n = (size_t)(N > 0 ? N : 1);
    return 2 * ((n * 4 + 255) & ~(size_t)255) + ((n / SCAN_TILE + 4) * 4 + 255) + n * 16 + 1024;
}
cudaError_t sgbk_fix_assignments(const int32_t* d_state, const int32_t* d_A, const int32_t* d_B, const int32_t* d_alleleA,
                                 const int32_t* d_alleleB, const uint8_t* d_seqhead, long long N, double loherror, int lik_bb, double rho,
                                 int32_t* d_A1, int32_t* d_B1, int32_t* d_A2, int32_t* d_B2, void* d_scratch, cudaStream_t st)
{
    if (N <= 0) return cudaSuccess;
    const double bb_s = (lik_bb && rho > 0.0) ? (1.0 - rho) / rho : 0.0;
    Carver cv{(char*)d_scratch, 0};
    unsigned* flag = cv.take<unsigned>((size_t)N);
    unsigned* rid = cv.take<unsigned>((size_t)N);
    unsigned* tsum = cv.take<unsigned>((size_t)N / SCAN_TILE + 4);
    unsigned long long* tot = cv.take<unsigned long long>(2 * (size_t)N);
    k_fix_pass1<<<grid_for(N), 256, 0, st>>>(d_state, d_A, d_B, d_alleleA, d_alleleB, d_seqhead, N, loherror, lik_bb, rho, bb_s, d_A1, d_B1);
    k_fix_runheads<<<grid_for(N), 256, 0, st>>>(d_A1, d_B1, d_seqhead, N, lik_bb, flag);
    cudaError_t e = scan_u32(flag, rid, N, tsum, nullptr, st);
    if (e != cudaSuccess) return e;
    k_fix_runids<<<grid_for(N), 256, 0, st>>>(flag, rid, N, tot);
    k_fix_runtotals<<<grid_for(N), 256, 0, st>>>(d_alleleA, d_alleleB, rid, N, tot);
    k_fix_pass2<<<grid_for(N), 256, 0, st>>>(d_state, d_A1, d_B1, rid, tot, N, loherror, lik_bb, rho, bb_s, d_A2, d_B2);
    return cudaGetLastError();
}
