// scan.cuh -- persistent-tile scheduling and a CTA-wide decoupled look-back.
//
// Variable-length output (hits per query, characters per view) needs an exclusive prefix
// over tiles.  All tiles of a batch are processed by ONE kernel: a tile publishes its
// aggregate, then the whole CTA (not one warp) inspects up to THREADS predecessors at
// once, so a batch of ~1000 simultaneously running tiles resolves in <= 4 L2 round trips
// instead of a 30-deep chain of 32-wide windows.  Status words carry the launch epoch,
// so no clearing pass runs between launches.
#pragma once
#include "common.cuh"

namespace hlsmap {

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int THREADS>
struct TileScan {
    static constexpr int NW = THREADS / 32;
    struct Smem {
        int tile;
        int warp_first[NW];
        unsigned long long warp_sum[NW];
    };

    // Called by every CTA before its first ticket.  The launches on a scratch set are numbered by the
    // host (seq = 0, 1, 2, ...; its low bit also picks the launch's ticket counter) and take the set
    // strictly in that order: `turn` is the number of launches that have finished with it, and a
    // launch's CTAs wait here until it is theirs.  (An earlier build handed the set to whichever
    // waiting launch won a compare-and-swap on an owner word: with several small launches in flight
    // under programmatic dependent launch, the launch after next could take the set first, draw from
    // a ticket counter nobody had reset yet, find no tiles and leave its batch untouched.)
    // Every CTA that calls this must call leave_set() before it exits.
    __device__ static void acquire_set(ScanCounters *ctr, unsigned seq)
    {
        if (threadIdx.x == 0) {
            while (*reinterpret_cast<volatile unsigned *>(&ctr->turn) != seq) __nanosleep(256);
            __threadfence();
        }
        __syncthreads();
    }

    // Tickets hand out tiles in order, so every predecessor of a tile is already owned
    // by a running CTA (forward progress of the look-back does not depend on blockIdx).
    __device__ static int next_tile(Smem &s, ScanCounters *ctr, unsigned tslot)
    {
        __syncthreads();  // previous tile fully consumed (shared buffers, s.tile)
        if (threadIdx.x == 0) s.tile = (int)atomicAdd(&ctr->ticket[tslot & 1u], 1u);
        __syncthreads();
        return s.tile;
    }

    // A batch may not produce more than INT32_MAX entries (the host reports OVERFLOW), so
    // published values saturate far above that instead of wrapping the 40-bit field.
    static constexpr unsigned long long kSat = 1ull << 33;

    // Step 1: make this tile's aggregate visible to later tiles (tile 0 publishes a prefix).
    // `base` = entries produced by earlier batches of the same logical call (chunked host
    // pipeline): it is the prefix of the virtual tile before tile 0.
    __device__ static void publish_aggregate(unsigned long long *status, unsigned epoch, int tile,
                                             unsigned long long aggregate,
                                             unsigned long long base = 0)
    {
        aggregate = aggregate < kSat ? aggregate : kSat;
        if (threadIdx.x == 0)
            st_relaxed_u64(&status[tile],
                           tile == 0 ? make_status(epoch, kStatePrefix, base + aggregate)
                                     : make_status(epoch, kStateAggregate, aggregate));
    }

    // Step 2: the sum of the aggregates of all earlier tiles.  Called by every thread of the
    // CTA after publish_aggregate(); work that does not need the prefix can sit in between.
    // Every step inspects SLOTS * THREADS predecessors at once (independent loads in flight):
    // the nearest tile that already knows its own prefix ends the walk.
    // Issues the loads of the first look-back window of `tile` early (e.g. before a search
    // phase) so their L2 round trip is hidden; hand `pre` to wait_prefix() later.
    template <int SLOTS>
    __device__ static void prefetch_window(const unsigned long long *status, int tile,
                                           unsigned long long (&pre)[SLOTS])
    {
#pragma unroll
        for (int j = 0; j < SLOTS; j++) {
            const int idx = tile - 1 - ((int)threadIdx.x + j * THREADS);
            pre[j] = idx >= 0 ? ld_relaxed_u64(&status[idx]) : 0ull;
        }
    }

    template <int SLOTS = 1>
    __device__ static unsigned long long wait_prefix(Smem &s, unsigned long long *status,
                                                     unsigned epoch, int tile,
                                                     unsigned long long aggregate,
                                                     const unsigned long long *pre = nullptr,
                                                     unsigned long long base = 0)
    {
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        if (tile == 0) return base;
        aggregate = aggregate < kSat ? aggregate : kSat;
        unsigned long long running = 0;
        int top = tile - 1;
        for (;;) {
            unsigned long long w[SLOTS];
#pragma unroll
            for (int j = 0; j < SLOTS; j++) {      // distance d = tid + j*THREADS behind `top`
                const int idx = top - (tid + j * THREADS);
                if (pre != nullptr && top == tile - 1) w[j] = pre[j];   // loaded ahead of time
                else w[j] = idx >= 0 ? ld_relaxed_u64(&status[idx]) : 0ull;
            }
            int my_first = 0x7fffffff;             // nearest distance holding a full prefix
            unsigned long long val[SLOTS];
#pragma unroll
            for (int j = 0; j < SLOTS; j++) {
                const int d = tid + j * THREADS;
                const int idx = top - d;
                bool is_prefix = true;             // idx < 0: the virtual tile -1 holds prefix 0
                val[j] = 0;
                if (idx >= 0) {
                    while ((unsigned)(w[j] >> 42) != (epoch & kEpochMask)) {
                        __nanosleep(128);          // not published yet: poll politely
                        w[j] = ld_relaxed_u64(&status[idx]);
                    }
                    val[j] = w[j] & ((1ull << kValueBits) - 1ull);
                    is_prefix = ((w[j] >> kValueBits) & 3ull) == kStatePrefix;
                }
                if (is_prefix) my_first = min(my_first, d);
            }
#pragma unroll
            for (int dd = 16; dd > 0; dd >>= 1) my_first = min(my_first, __shfl_xor_sync(0xffffffffu, my_first, dd));
            __syncthreads();  // s.warp_first / s.warp_sum free for reuse
            if (lane == 0) s.warp_first[warp] = my_first;
            __syncthreads();
            int first = 0x7fffffff;
#pragma unroll
            for (int ww = 0; ww < NW; ww++) first = min(first, s.warp_first[ww]);
            unsigned long long c = 0;
#pragma unroll
            for (int j = 0; j < SLOTS; j++) c += (tid + j * THREADS <= first) ? val[j] : 0ull;
#pragma unroll
            for (int dd = 16; dd > 0; dd >>= 1) c += __shfl_xor_sync(0xffffffffu, c, dd);
            if (lane == 0) s.warp_sum[warp] = c;
            __syncthreads();
#pragma unroll
            for (int ww = 0; ww < NW; ww++) running += s.warp_sum[ww];
            if (first != 0x7fffffff) break;
            top -= SLOTS * THREADS;
        }
        running = running < kSat ? running : kSat;
        if (tid == 0)
            st_relaxed_u64(&status[tile], make_status(epoch, kStatePrefix, running + aggregate));
        return running;
    }

    // The same look-back done by ONE warp, without any CTA barrier: 32 predecessors per step, the
    // walk ends at the nearest tile that knows its prefix (in the search kernel's pipeline, where a
    // tile asks one stage after it published, that is 2 steps on average and 1 for half the tiles).
    // `pre` = status[tile - 1 - lane], loaded ahead of time by this warp (ignored for tile 0).
    // Every lane gets the result.
    __device__ static unsigned long long warp_prefix(unsigned long long *status, unsigned epoch, int tile,
                                                     unsigned long long aggregate, unsigned long long pre,
                                                     unsigned long long base, int *diag = nullptr)
    {
        if (tile == 0) return base;                  // tile 0 published its own prefix
        int n_windows = 0, n_polls = 0;
        const int lane = threadIdx.x & 31;
        const unsigned e = epoch & kEpochMask;
        const unsigned long long vmask = (1ull << kValueBits) - 1ull;
        aggregate = aggregate < kSat ? aggregate : kSat;
        unsigned long long running = 0;
        int top = tile - 1;
        unsigned long long w = pre;
        for (;;) {
            const int idx = top - lane;
            int n;                                   // lanes 0..n-1 are summed, then lane n if it is a prefix
            bool ends;
            for (;;) {
                const bool pub = idx < 0 || (unsigned)(w >> 42) == e;
                const bool isp = idx < 0 || (pub && ((w >> kValueBits) & 3ull) == kStatePrefix);
                const unsigned m_pub = __ballot_sync(0xffffffffu, pub);
                const unsigned m_pre = __ballot_sync(0xffffffffu, isp);
                const int first_unpub = m_pub == 0xffffffffu ? 32 : __ffs(~m_pub) - 1;
                const int first_pre = m_pre == 0u ? 32 : __ffs(m_pre) - 1;
                if (first_pre < first_unpub) { n = first_pre; ends = true; break; }
                if (first_unpub == 32) { n = 31; ends = false; break; }
                __nanosleep(64);                     // a nearer tile has not published yet
                n_polls++;
                if (!pub) w = ld_relaxed_u64(&status[idx]);
            }
            n_windows++;
            unsigned long long c = (lane <= n && idx >= 0) ? (w & vmask) : 0ull;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
            running += c;
            if (ends) break;
            top -= 32;
            const int nidx = top - lane;
            w = nidx >= 0 ? ld_relaxed_u64(&status[nidx]) : 0ull;
        }
        running = running < kSat ? running : kSat;
        if (lane == 0) st_relaxed_u64(&status[tile], make_status(epoch, kStatePrefix, running + aggregate));
        if (diag) *diag = n_windows | (n_polls << 8);
        return running;
    }

    __device__ static unsigned long long exclusive_prefix(Smem &s, unsigned long long *status,
                                                          unsigned epoch, int tile,
                                                          unsigned long long aggregate,
                                                          unsigned long long base = 0)
    {
        publish_aggregate(status, epoch, tile, aggregate, base);
        return wait_prefix(s, status, epoch, tile, aggregate, nullptr, base);
    }

    // Call once per finished tile, after its global writes.  Returns true in thread 0 of
    // exactly one CTA: the one that finished the batch's last outstanding tile.
    __device__ static bool tile_done(ScanCounters *ctr, int n_tiles)
    {
        __syncthreads();
        if (threadIdx.x != 0) return false;
        __threadfence();
        const unsigned old = atomicAdd(&ctr->done, 1u);
        if (old != (unsigned)(n_tiles - 1)) return false;
        __threadfence();
        return true;
    }

    // Run by the thread tile_done() elected: hand the counters to the host-visible
    // result and re-arm the scratch header for the next launch.  The set itself stays this
    // launch's until its last CTA has left (leave_set): a CTA that is dispatched only after
    // the batch's last tile was finished still finds turn == its launch's number, draws a ticket
    // past the end and leaves; were the set passed on here, that late CTA would wait for a turn
    // that has gone by.
    __device__ static void finish_batch(ScanCounters *ctr, DeviceResult *result, unsigned tslot)
    {
        result->n_unparsed = (long long)atomicAdd(&ctr->n_unparsed, 0ull);
        result->n_missing = (long long)atomicAdd(&ctr->n_missing, 0ull);
        ctr->n_unparsed = 0ull;
        ctr->n_missing = 0ull;
        ctr->done = 0u;
        ctr->ticket[(tslot & 1u) ^ 1u] = 0u;       // untouched by this launch's late fetches
        __threadfence();
    }

    // Called by every CTA as the last thing it does.  The last CTA of the grid to get here
    // passes the set on to the next launch on it (turn = seq + 1).  A kernel that does not count
    // finished tiles (tile_done) passes `result`: the batch is then wrapped up here, where every
    // tile is known to be done because every CTA has left.
    __device__ static void leave_set(ScanCounters *ctr, unsigned seq, DeviceResult *result = nullptr)
    {
        __syncthreads();                            // this CTA's finish_batch(), if any, is done
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned old = atomicAdd(&ctr->exited, 1u);
            if (old == gridDim.x - 1u) {
                __threadfence();
                if (result != nullptr) finish_batch(ctr, result, seq);
                ctr->exited = 0u;
                __threadfence();
                atomicExch(&ctr->turn, seq + 1u);
            }
        }
    }
};

}  // namespace hlsmap
