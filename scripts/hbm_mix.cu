// hbm_mix.cu -- what HBM3e on this B200 sustains for the traffic MIX of the list form:
// read-only, write-only, copy, and "24 B read + 153 B written per query" streams.
// The copy figure in MEASURED_PEAKS.json is half reads, half writes; getSegmentList on the
// bench workload is 85 % writes, so its attainable ceiling is the write-heavy figure below.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/hbm_mix scripts/hbm_mix.cu && /tmp/hbm_mix
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_write(double *o, size_t n, double v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * st < n; i += 4 * st) {
        __stcs(o + i, v); __stcs(o + i + st, v); __stcs(o + i + 2 * st, v); __stcs(o + i + 3 * st, v);
    }
    for (; i < n; i += st) __stcs(o + i, v);
}
__global__ void k_write12(double *o, int *o2, size_t n, double v)   // 8 B + 4 B per slot, like the hit columns
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * st < n; i += 4 * st) {
#pragma unroll
        for (int j = 0; j < 4; j++) { __stcs(o + i + j * st, v); __stcs(o2 + i + j * st, (int)i); }
    }
    for (; i < n; i += st) { __stcs(o + i, v); __stcs(o2 + i, (int)i); }
}
__global__ void k_read(const double *a, size_t n, double *sink)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    double s = 0;
    for (; i + 3 * st < n; i += 4 * st)
        s += __ldcs(a + i) + __ldcs(a + i + st) + __ldcs(a + i + 2 * st) + __ldcs(a + i + 3 * st);
    for (; i < n; i += st) s += __ldcs(a + i);
    if (s == 1.2345e-300) *sink = s;
}
__global__ void k_copy(const double *a, double *o, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * st < n; i += 4 * st) {
        double x0 = __ldcs(a + i), x1 = __ldcs(a + i + st), x2 = __ldcs(a + i + 2 * st), x3 = __ldcs(a + i + 3 * st);
        __stcs(o + i, x0); __stcs(o + i + st, x1); __stcs(o + i + 2 * st, x2); __stcs(o + i + 3 * st, x3);
    }
    for (; i < n; i += st) __stcs(o + i, __ldcs(a + i));
}
// per "query": 3 doubles read, `h` doubles + h ints written (h = 12.77 on the bench workload -> 13)
__global__ void k_mix(const double *a, double *o, int *o2, size_t nq, int h)
{
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; q < nq; q += st) {
        double x = __ldcs(a + q) + __ldcs(a + nq + q) + __ldcs(a + 2 * nq + q);
        for (int j = 0; j < h; j++) { __stcs(o + (size_t)j * nq + q, x); __stcs(o2 + (size_t)j * nq + q, (int)q); }
    }
}

// the same 8 B + 4 B columns, written by bulk copies out of shared memory (TMA, no LSU stores):
// every CTA stages CH slots and hands them to the copy engine, two buffers in flight
__global__ void k_write12_bulk(double *o, int *o2, size_t n, double v)
{
    constexpr int CH = 1024;
    __shared__ __align__(128) double sd[2][CH];
    __shared__ __align__(128) int si[2][CH];
    const size_t n_chunks = n / CH;
    int b = 0;
    for (size_t c = blockIdx.x; c < n_chunks; c += gridDim.x, b ^= 1) {
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // buffer b is free again
        __syncthreads();
        for (int i = threadIdx.x; i < CH; i += blockDim.x) { sd[b][i] = v; si[b][i] = (int)i; }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o + c * CH),
                         "r"((unsigned)__cvta_generic_to_shared(&sd[b][0])), "r"(CH * 8) : "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o2 + c * CH),
                         "r"((unsigned)__cvta_generic_to_shared(&si[b][0])), "r"(CH * 4) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// The list form's traffic in the list form's SHAPE, without any of its arithmetic: persistent CTAs take
// 512-query tiles in order; a tile reads its queries (4 + 8 + 8 B each, three contiguous runs), writes
// 512 offsets and `hpt` hits as one contiguous run of doubles and one of ints at hpt * tile.
__global__ void k_mix_tiled(const int *lv, const double *bg, const double *du, int *off, double *tm, int *sn,
                            size_t n_tiles, int hpt, unsigned *ticket)
{
    __shared__ unsigned s_tile;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const size_t tile = s_tile;
        if (tile >= n_tiles) break;
        const size_t q0 = tile * 512 + threadIdx.x * 2;
        const int2 l = __ldcs(reinterpret_cast<const int2 *>(lv + q0));
        const double2 b = __ldcs(reinterpret_cast<const double2 *>(bg + q0));
        const double2 d = __ldcs(reinterpret_cast<const double2 *>(du + q0));
        const double x = b.x + d.x + b.y + d.y + l.x + l.y;
        __stcs(reinterpret_cast<int2 *>(off + q0), make_int2((int)q0, (int)q0 + 1));
        double *t = tm + tile * hpt;
        int *s = sn + tile * hpt;
        for (int i = threadIdx.x; i < hpt; i += blockDim.x) {
            __stcs(t + i, x);
            __stcs(s + i, i);
        }
    }
}

// The same tiles, their hits written by bulk copies out of shared memory (1024 slots per copy pair,
// two buffers in flight) instead of by per-thread stores.  hpt must be a multiple of 4 (alignment).
__global__ void k_mix_tiled_bulk(const int *lv, const double *bg, const double *du, int *off, double *tm, int *sn,
                                 size_t n_tiles, int hpt, unsigned *ticket)
{
    constexpr int CH = 1024;
    __shared__ __align__(128) double sd[2][CH];
    __shared__ __align__(128) int si[2][CH];
    __shared__ unsigned s_tile;
    int b = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const size_t tile = s_tile;
        if (tile >= n_tiles) break;
        const size_t q0 = tile * 512 + threadIdx.x * 2;
        const int2 l = __ldcs(reinterpret_cast<const int2 *>(lv + q0));
        const double2 bb = __ldcs(reinterpret_cast<const double2 *>(bg + q0));
        const double2 d = __ldcs(reinterpret_cast<const double2 *>(du + q0));
        const double x = bb.x + d.x + bb.y + d.y + l.x + l.y;
        __stcs(reinterpret_cast<int2 *>(off + q0), make_int2((int)q0, (int)q0 + 1));
        double *t = tm + tile * hpt;
        int *s = sn + tile * hpt;
        for (int c0 = 0; c0 < hpt; c0 += CH, b ^= 1) {
            const int n = hpt - c0 < CH ? hpt - c0 : CH;
            if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncthreads();
            for (int i = threadIdx.x; i < n; i += blockDim.x) { sd[b][i] = x; si[b][i] = i; }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(t + c0),
                             "r"((unsigned)__cvta_generic_to_shared(&sd[b][0])), "r"(n * 8) : "memory");
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(s + c0),
                             "r"((unsigned)__cvta_generic_to_shared(&si[b][0])), "r"(n * 4) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ... and with every WARP staging 128 slots of its own and issuing its own pair of small bulk copies
// (1 KB + 512 B), no CTA barrier in the loop: the granularity a kernel can use whose warps run free.
__global__ void k_mix_tiled_bulk_warp(const int *lv, const double *bg, const double *du, int *off, double *tm, int *sn,
                                      size_t n_tiles, int hpt, unsigned *ticket)
{
    constexpr int CH = 128;
    __shared__ __align__(128) double sd[8][2][CH];
    __shared__ __align__(128) int si[8][2][CH];
    __shared__ unsigned s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int b = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const size_t tile = s_tile;
        if (tile >= n_tiles) break;
        const size_t q0 = tile * 512 + threadIdx.x * 2;
        const int2 l = __ldcs(reinterpret_cast<const int2 *>(lv + q0));
        const double2 bb = __ldcs(reinterpret_cast<const double2 *>(bg + q0));
        const double2 d = __ldcs(reinterpret_cast<const double2 *>(du + q0));
        const double x = bb.x + d.x + bb.y + d.y + l.x + l.y;
        __stcs(reinterpret_cast<int2 *>(off + q0), make_int2((int)q0, (int)q0 + 1));
        double *t = tm + tile * hpt;
        int *s = sn + tile * hpt;
        for (int c0 = warp * CH; c0 + CH <= hpt; c0 += 8 * CH, b ^= 1) {      // (a ragged end is left out: same bytes +- 1 %)
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; j++) { sd[warp][b][lane + 32 * j] = x; si[warp][b][lane + 32 * j] = c0 + j; }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(t + c0),
                             "r"((unsigned)__cvta_generic_to_shared(&sd[warp][b][0])), "r"(CH * 8) : "memory");
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(s + c0),
                             "r"((unsigned)__cvta_generic_to_shared(&si[warp][b][0])), "r"(CH * 4) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class F>
static float timed(F f, int reps)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f(); f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    const size_t n = (size_t)1 << 28;   // 2 GiB of doubles per buffer: far beyond the 126 MB L2
    double *a, *o, *sink; int *o2;
    cudaMalloc(&a, n * 8); cudaMalloc(&o, n * 8); cudaMalloc(&o2, n * 4); cudaMalloc(&sink, 8);
    cudaMemset(a, 0, n * 8); cudaMemset(o, 0, n * 8); cudaMemset(o2, 0, n * 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int cps : {1, 2, 4}) {      // how fast can ONE or TWO 256-thread CTAs per SM store?
        const int grid = sms * cps, th = 256;
        float w = timed([&] { k_write<<<grid, th>>>(o, n / 4, 1.0); }, 5);
        float w12 = timed([&] { k_write12<<<grid, th>>>(o, o2, n / 4, 1.0); }, 5);
        float wb = timed([&] { k_write12_bulk<<<grid, th>>>(o, o2, n / 4, 1.0); }, 5);
        printf("{\"ctas_per_sm\": %d, \"write_GBps\": %.0f, \"write_8plus4_GBps\": %.0f, \"write_8plus4_bulk_GBps\": %.0f}\n",
               cps, n / 4 * 8 / w / 1e6, n / 4 * 12 / w12 / 1e6, n / 4 * 12 / wb / 1e6);
    }
    {   // 2^22 queries in 512-query tiles, 6537 hits per tile (12.77 per query): 4 CTAs per SM as the kernel runs
        const size_t nq = (size_t)1 << 22, n_tiles = nq / 512;
        const int hpt = 6537;
        unsigned *ticket;
        cudaMalloc(&ticket, 4);
        int *lv = reinterpret_cast<int *>(a);
        double *bg = a + nq, *du = a + 2 * nq;
        for (int cps : {3, 4, 8}) {
            float m = timed([&] {
                cudaMemsetAsync(ticket, 0, 4);
                k_mix_tiled<<<sms * cps, 256>>>(lv, bg, du, o2 + (n - nq - 16), o, o2, n_tiles, hpt, ticket);
            }, 5);
            const int hb = 6536;
            float mb = timed([&] {
                cudaMemsetAsync(ticket, 0, 4);
                k_mix_tiled_bulk<<<sms * cps, 256>>>(lv, bg, du, o2 + (n - nq - 16), o, o2, n_tiles, hb, ticket);
            }, 5);
            float mw = timed([&] {
                cudaMemsetAsync(ticket, 0, 4);
                k_mix_tiled_bulk_warp<<<sms * cps, 256>>>(lv, bg, du, o2 + (n - nq - 16), o, o2, n_tiles, hb, ticket);
            }, 5);
            // the same per-thread stores with every tile's run starting on a line boundary (6528 = 51 * 128
            // slots per tile), and on a 64-byte boundary of the 8-byte column (6536)
            for (int h2 : {6528, 6536}) {
                float ma = timed([&] {
                    cudaMemsetAsync(ticket, 0, 4);
                    k_mix_tiled<<<sms * cps, 256>>>(lv, bg, du, o2 + (n - nq - 16), o, o2, n_tiles, h2, ticket);
                }, 5);
                printf("{\"ctas_per_sm\": %d, \"hits_per_tile\": %d, \"mix_tiled_per_thread_stores_GBps\": %.0f}\n", cps, h2,
                       (nq * 28.0 + n_tiles * 12.0 * h2) / ma / 1e6);
            }
            printf("{\"ctas_per_sm\": %d, \"mix_tiled_like_the_list_form_GBps\": %.0f, \"same_with_bulk_stores_GBps\": %.0f, "
                   "\"same_with_per_warp_bulk_stores_GBps\": %.0f}\n", cps,
                   (nq * 28.0 + n_tiles * 12.0 * hpt) / m / 1e6, (nq * 28.0 + n_tiles * 12.0 * hb) / mb / 1e6,
                   (nq * 28.0 + n_tiles * 12.0 * (hb / 128 * 128)) / mw / 1e6);
        }
    }
    for (int cps : {4, 8, 16}) {
        const int grid = sms * cps, th = 256;
        float w = timed([&] { k_write<<<grid, th>>>(o, n, 1.0); }, 5);
        float w12 = timed([&] { k_write12<<<grid, th>>>(o, o2, n, 1.0); }, 5);
        float r = timed([&] { k_read<<<grid, th>>>(a, n, sink); }, 5);
        float c = timed([&] { k_copy<<<grid, th>>>(a, o, n); }, 5);
        const size_t nq = (size_t)1 << 24; const int h = 13;
        float m = timed([&] { k_mix<<<grid, th>>>(a, o, o2, nq, h); }, 5);
        printf("{\"ctas_per_sm\": %d, \"write_GBps\": %.0f, \"write_8plus4_GBps\": %.0f, \"read_GBps\": %.0f, "
               "\"copy_GBps\": %.0f, \"mix_24r_156w_GBps\": %.0f}\n",
               cps, n * 8 / w / 1e6, n * 12 / w12 / 1e6, n * 8 / r / 1e6, 2.0 * n * 8 / c / 1e6,
               nq * (24.0 + 12.0 * h) / m / 1e6);
    }
    return 0;
}
