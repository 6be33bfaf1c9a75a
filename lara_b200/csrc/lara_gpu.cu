// lara_gpu.cu -- C ABI (include/lara_b200.h) and host orchestration of the GPU hot path.
//
// Drop-in seam: lara::solve(results, store, params)  (src/subgradient_solver.hpp:110, src/lara.cpp:57).
// The reference keeps `threads` solver slots alive and refills them from a longest-first queue
// (src/subgradient_solver.hpp:112-120, 286-325).  Here the queue becomes length-sorted BATCHES of pairs
// that iterate in lock step; each iteration is one DP launch per kernel class present in the batch (band
// shape (G, Q) or rows per lane R, side by side on streams) plus one post-processing launch.  Several
// workers per device pull batches from the device's queue and run them concurrently.  Pairs are independent,
// so several GPUs simply take interleaved shares of the sorted pair list (one host thread per device, no
// collective).
#include "lara_gpu.cuh"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <tuple>
#include <string>
#include <thread>
#include <vector>

using namespace lara;

namespace
{

thread_local std::string g_createErr;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_));          \
    } while (0)

struct OrientedPair
{
    int32_t first, second;      // first = longer sequence (ties: lower index), like solver :117-120
    int32_t la, lb;
    int64_t request;            // position in the caller's pair list
};

// One host worker of a device: its own stream, events and live-pair counter.  Several workers run batches
// concurrently so that the issue-bound DP kernel of one batch overlaps the latency-bound post kernel of another
// and finished pairs of one batch leave room for the blocks of the others.
struct Worker
{
    cudaStream_t stream = nullptr;
    int32_t * hFinished = nullptr;          // pinned: two windows of the device's finished list (count + entries)
    size_t hFinishedCap = 0;
    cudaEvent_t evBlock[2] = {nullptr, nullptr};   // end of the last two blocks of iteration steps
    int32_t * hTable = nullptr;             // pinned staging buffer of the slot table
    size_t hTableCap = 0;
    cudaEvent_t evTable = nullptr;
    std::vector<cudaEvent_t> events;
    cudaEvent_t evDone = nullptr;
    // side streams: the DP launches of the smaller kernel classes of an iteration run beside the largest one
    std::vector<cudaStream_t> aux;
    std::vector<cudaEvent_t> auxJoin;
    cudaEvent_t evFork = nullptr;
    // stream and event of the set-up of refill cohorts (made once: creating a stream takes the driver's lock)
    cudaStream_t side = nullptr;
    cudaEvent_t evSide = nullptr;
    // slab of the batch arrays: grown between batches only, so that the iteration loop never waits for the allocator
    int32_t * slab = nullptr;
    int64_t slabInts = 0;
    lara_stats stats{};
    std::string err;
};

// Density of a sample of a device's pairs (edges, interactions and band bytes per matrix cell), measured by running
// the set-up stages on it; sizes the batches (batch_limit).
struct Pilot
{
    double cells = 0, edges = 0, inters = 0, dpBytes = 0;
    int bandSlack = 0;      // largest (diagonals the active edges span) - (la - lb + 1) of the sample: what a band needs beyond
                            // the diagonals between the matrix's two corners
    bool band = false;      // the sample ran the band kernels
};

// All device-side resources of one GPU.
struct Device
{
    int id = 0;
    int smCount = 148;
    int smemOptin = 0;                      // cudaDevAttrMaxSharedMemoryPerBlockOptin
    cudaStream_t stream = nullptr;          // == workers[0].stream
    std::vector<Worker> workers;
    std::mutex queueMutex;
    size_t nextBegin = 0;                   // next unassigned pair of `pairs`
    DevInputs in{};
    void * dSeq = nullptr, * dSeqOff = nullptr, * dBpPartner = nullptr, * dBpWeight = nullptr, * dBpOff = nullptr;
    // capacities in bytes of the input and output arrays: a call that fits what the call before allocated makes no
    // cudaMalloc / cudaFree (both take a driver lock that other processes on the box hold for 0.2-0.7 s at times:
    // profiles/README.md r2v)
    size_t capSeq = 0, capSeqOff = 0, capBpPartner = 0, capBpWeight = 0, capBpOff = 0, capOutHdr = 0, capOutLines = 0, capOutOff = 0;
    int32_t * negBuf = nullptr;
    std::vector<cudaEvent_t> events;        // two scratch events for copies / microbenchmarks
    cudaEvent_t evRun0 = nullptr, evRun1 = nullptr;     // span of one run_device on the stream
    // share of the prepared pair list and its device-resident outputs
    std::vector<OrientedPair> pairs;
    Pilot pilot;                            // densities of a sample of `pairs`
    PairOutHdr * dOutHdr = nullptr;
    int32_t * dOutLines = nullptr;
    int64_t * dOutOff = nullptr;
    std::vector<int64_t> outOff;
    int64_t outLinesTotal = 0;
    lara_stats stats{};
    std::string err;
    bool poolHeld = false;                  // counted in g_poolUsers
};

} // namespace

struct lara_ctx
{
    lara_params P{};
    DevParams dp{};
    std::string err;
    std::deque<Device> devs;                // deque: Device holds a mutex and is not movable
    int nWorkers = 2;                       // concurrent batches per device (env LARA_B200_WORKERS)
    size_t batchPairsDefault = 56832;       // pairs per batch (env LARA_B200_BATCH): 148 SMs x 12 warps x 32 one-thread pairs
    bool bandMode = true;                   // band DP kernels where the active edges allow it (env LARA_B200_BAND=0: off)
    int bandMargin = 1;                     // diagonals added on both sides of the active range (env LARA_B200_MARGIN)
    int bandMaxW = 1024;                    // widest band; beyond it the full-matrix kernel runs (env LARA_B200_BANDMAXW)
    int bandMinWarps = 600;                 // a batch whose fill has fewer warps than this with blocks of 16 diagonals runs
                                            // with blocks of 8 and twice the lanes per pair (env LARA_B200_BANDMINWARPS)
    int bandClassMinWarps = 64;             // ... and a width class of fewer warps than this too (env LARA_B200_BANDCLASSMINWARPS)
    int band24MinWarps = 64;                // the 24-diagonal shape needs this many warps of its own (env LARA_B200_BAND24MINWARPS)
    bool band24 = true;                     // 24-diagonal one-thread-per-pair shape (env LARA_B200_BAND24=0: off)
    int bandForceQ = 0;                     // 4 / 16: fixed block size (env LARA_B200_BANDQ)
    int verbose = 0;                        // env LARA_B200_VERBOSE=1: host-side stage times of every batch on stderr
    int postSmemCand = 256;                 // matching candidates sorted in shared memory (env LARA_B200_SMEMCAND)
    // host copy of the inputs (for pair orientation and batch sizing)
    int32_t nSeq = 0;
    std::vector<int64_t> seqOff, bpOff;
    std::vector<int32_t> nbp;               // base pairs per sequence
    int64_t nPrepared = 0;
    bool prepared = false, ran = false;
    lara_stats stats{};
    size_t batchPairsCap = 0;               // test hook: max pairs per batch, 0 = automatic (env LARA_B200_BATCHCAP)
    bool postInOrder = true;                // post kernel over the pairs in admission order instead of slot order (env LARA_B200_POSTORDER=0: slot order)
    bool splitLaunch = false;               // env LARA_B200_SPLIT=1: band classes of >= bandSplitWarps warps (except the largest)
    int bandSplitWarps = 256;               // run as launches of their own (LARA_B200_SPLITWARPS)
    int rowsMode = 0;                       // form of the row-by-row kernel (kernels_rows.cu): 0 affine go <= ge, 1 linear gap
                                            // costs (gap open == gap extend in the scalar build's arithmetic), 2 affine go > ge
    bool forceRows = false;                 // every pair without a band runs the row-by-row kernel (env LARA_B200_ROWS=1; tests)
    int refillMode = 1;                     // slot refill: 0 = set by set, 1 = as slots free up while the queue is long,
                                            // 2 = always as slots free up (env LARA_B200_REFILL)
    double memBudget = 0;                   // test hook: bytes the workers of a device may hold together, 0 = automatic
                                            // (env LARA_B200_MEMBUDGET, in MB)
    double refillFirstShare = 0.8;          // share of a worker's memory budget the first cohort takes when slots will be
                                            // refilled as they free up (env LARA_B200_REFILLSHARE)
    size_t workerMinPairs = 2048;           // a device runs one worker per this many pairs, up to nWorkers (env LARA_B200_WORKERPAIRS)
};

namespace
{

// Page-locked host blocks for the result arrays (device-to-host copies at full PCIe / C2C speed).  Pinning is slow
// (~0.2 ms per MB), so freed blocks are kept for the next result set of similar size.
struct PinnedCache
{
    std::mutex m;
    std::vector<std::pair<void *, size_t>> blocks;
    void * acquire(size_t bytes, size_t & got)
    {
        {
            std::lock_guard<std::mutex> lock(m);
            for (size_t i = 0; i < blocks.size(); ++i)
                if (blocks[i].second >= bytes && blocks[i].second <= 2 * bytes + (1u << 20))
                {
                    void * p = blocks[i].first;
                    got = blocks[i].second;
                    blocks.erase(blocks.begin() + (long)i);
                    return p;
                }
        }
        void * p = nullptr;
        if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        got = bytes;
        return p;
    }
    void release(void * p, size_t bytes)
    {
        std::lock_guard<std::mutex> lock(m);
        blocks.emplace_back(p, bytes);
        while (blocks.size() > 2) { cudaFreeHost(blocks.front().first); blocks.erase(blocks.begin()); }
    }
};
PinnedCache g_pinned;

struct ResultOwner
{
    std::vector<lara_pair_result> pairs;
    // one block: posLow | posHigh | weight (not initialised: the device-to-host copies write every element);
    // page-locked when possible, else plain memory
    void * block = nullptr;
    size_t blockBytes = 0;
    bool pinned = false;
    int32_t * posLow = nullptr, * posHigh = nullptr;
    uint32_t * weight = nullptr;
    int64_t nLines = 0;
    void allocate(int64_t n)
    {
        size_t const each = (size_t)std::max<int64_t>(n, 1);
        block = g_pinned.acquire(3 * each * 4, blockBytes);
        pinned = block != nullptr;
        if (!block) { block = ::operator new(3 * each * 4); blockBytes = 3 * each * 4; }
        posLow = static_cast<int32_t *>(block);
        posHigh = posLow + each;
        weight = reinterpret_cast<uint32_t *>(posHigh + each);
    }
    ~ResultOwner()
    {
        if (block && pinned) g_pinned.release(block, blockBytes);
        else if (block) ::operator delete(block);
    }
};

DevParams make_dev_params(lara_params const & P)
{
    DevParams d{};
    d.numIterations = P.num_iterations;
    d.maxNondecreasing = P.max_nondecreasing;
    d.lookahead = P.matching;
    d.stepMode = P.step_mode;
    d.stepSizeFactor = P.step_size_factor;
    d.balance = P.balance;
    d.sequenceScale = P.sequence_scale;
    d.subopt = (int32_t)(P.suboptimal_diff * LARA_FACTOR2INT);      // src/lagrange.hpp:251
    d.go = P.gap_open;
    d.ge = P.gap_extend;
    for (int i = 0; i < 25; ++i) d.tab[i] = P.score_matrix[i];
    d.libMin = P.lib_score_min;
    d.libMax = P.lib_score_max;
    d.libLinear = P.lib_score_linear;
    return d;
}

// cudaMallocAsync that survives a full pool: the workers of a device size their sets from an ESTIMATE of what a pair will
// hold, and blocks freed on one stream are not at once usable from another.  On "out of memory" the device is drained
// (pending cudaFreeAsync calls of all workers take effect), the pool's cached blocks go back to the driver, and the call
// is repeated; while another worker still holds its set the wait goes on in steps of 10 ms, 20 times at most.
cudaError_t malloc_async_retry(void ** p, size_t bytes, cudaStream_t st)
{
    // test hook LARA_B200_FAILALLOC=k: every k-th call is told "out of memory" once before it is made
    char const * const fe = std::getenv("LARA_B200_FAILALLOC");
    int const failEvery = fe ? std::atoi(fe) : 0;
    static std::atomic<int> calls{0};
    bool const pretend = failEvery > 0 && (++calls % failEvery) == 0;
    cudaError_t e = pretend ? cudaErrorMemoryAllocation : cudaMallocAsync(p, bytes, st);
    for (int attempt = 0; e == cudaErrorMemoryAllocation && attempt < 20; ++attempt)
    {
        (void)cudaGetLastError();
        cudaError_t s2 = cudaDeviceSynchronize();
        if (s2 != cudaSuccess) return s2;
        int dev = 0;
        cudaMemPool_t pool = nullptr;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess)
            (void)cudaMemPoolTrimTo(pool, 0);
        e = cudaMallocAsync(p, bytes, st);
        if (e == cudaErrorMemoryAllocation) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    return e;
}

inline int64_t align4(int64_t x) { return (x + 3) & ~(int64_t)3; }

// Bump allocator over one device allocation; offsets are relative to the batch's base pointer (in ints).
struct Region
{
    int32_t * ptr = nullptr;
    int64_t size = 0, used = 0;
    int64_t baseDelta = 0;                  // (ptr - arenaBase) in ints
    int64_t take(int64_t nInts)
    {
        int64_t const o = used;
        used += align4(std::max<int64_t>(nInts, 1));
        return baseDelta + o;
    }
};

void device_free_inputs(Device & dv)
{
    cudaFree(dv.dSeq); cudaFree(dv.dSeqOff); cudaFree(dv.dBpPartner); cudaFree(dv.dBpWeight); cudaFree(dv.dBpOff);
    dv.dSeq = dv.dSeqOff = dv.dBpPartner = dv.dBpWeight = dv.dBpOff = nullptr;
    dv.capSeq = dv.capSeqOff = dv.capBpPartner = dv.capBpWeight = dv.capBpOff = 0;
}

void device_free_outputs(Device & dv)
{
    cudaFree(dv.dOutHdr); cudaFree(dv.dOutLines); cudaFree(dv.dOutOff);
    dv.dOutHdr = nullptr; dv.dOutLines = nullptr; dv.dOutOff = nullptr;
    dv.capOutHdr = dv.capOutLines = dv.capOutOff = 0;
}

// device array of at least `bytes`: kept when it is large enough, else freed and allocated anew
template <typename T>
void ensure_device(T *& p, size_t & cap, size_t bytes)
{
    bytes = std::max<size_t>(bytes, 1);
    if (p && cap >= bytes) return;
    if (p) { CK(cudaFree(p)); p = nullptr; cap = 0; }
    CK(cudaMalloc((void **)&p, bytes));
    cap = bytes;
}

double ev_ms(cudaEvent_t a, cudaEvent_t b)
{
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return (double)ms;
}

// Band geometry of one pair from the diagonal range of its active edges (kernels_band.cu): dLo <= 0 <= dHi,
// bandW = 32 .. 1024 diagonals (multiple of 32).  bandW = 0 selects the full-matrix kernel (band wider than 1024
// diagonals, or gap costs outside the band kernel's precondition go < ge).
void choose_band(lara_ctx const & ctx, PairDesc & d, bool allow24)
{
    d.bandW = 0; d.bandG = 0; d.bandQ = 0; d.bandSlot = 0; d.dLo = 0; d.dHi = -1;
    if (!ctx.bandMode || !(ctx.dp.go < ctx.dp.ge)) return;
    int lo = 0, hi = 0;
    if (d.E > 0 && d.dMin <= d.dMax)
    {
        lo = std::min(d.dMin - ctx.bandMargin, 0);
        hi = std::max(d.dMax + ctx.bandMargin, 0);
    }
    int const need = hi - lo + 1;
    int W = (allow24 && need <= 24) ? 24 : 32;
    while (W < need) W <<= 1;
    if (W > ctx.bandMaxW) return;
    // a band that needs the full-matrix kernel's work or more is not worth it (short sequences)
    if (d.lb <= 32 * kMaxR && (int64_t)W * (d.lb + 1) >= (int64_t)d.la * d.lb) return;
    int const slack = W - need;
    lo -= std::min(slack / 2, std::max(0, d.lb + lo));          // spare diagonals: both sides, none below -lb
    d.bandW = W; d.dLo = lo; d.dHi = lo + W - 1;
}

// Lanes per pair G and diagonals per block Q for the pairs of a batch that share the band width W.  One thread walks
// a block of Q diagonals down the rows as one dependent instruction stream, so the time of a launch is set by Q (measured
// per row: about 2,000 cycles at Q = 16, 900 at Q = 8, 1,240 at Q = 4 where the hand-off between lanes dominates) as
// long as the launch does not fill the machine.  batchWarps16 = warps the whole batch would have with blocks of 16: with
// at least bandMinWarps of them (about one per warp scheduler) the throughput form Q = 16 is used (G = W / 32, down to
// one thread per pair), else the low-latency form Q = 8 (G = W / 16).  Q = 4 only on request (LARA_B200_BANDQ).
void choose_band_shape(lara_ctx const & ctx, int W, int64_t batchWarps16, int64_t classWarps16, int & G, int & Q)
{
    if (W == 24) { G = 1; Q = 12; return; }                     // one thread per pair only
    int const g8 = W / 16, g4 = W / 8;
    // all shapes of a set share one launch, which lasts as long as its slowest class: a class of a few warps (the
    // handful of C3 pairs that need 64 diagonals) takes the low-latency form even in a big set
    bool const throughput = batchWarps16 >= ctx.bandMinWarps && classWarps16 >= ctx.bandClassMinWarps;
    Q = (g8 > 32 || throughput) ? 16 : 8;
    if (ctx.bandForceQ == 16 || (ctx.bandForceQ == 8 && g8 <= 32) || (ctx.bandForceQ == 4 && g4 <= 32)) Q = ctx.bandForceQ;
    G = W / (2 * Q);
}

// ---------------------------------------------------------------------------------------------------
// The resident set of one worker: the pairs that are iterating on the device right now.
//
// The reference keeps `threads` solver slots busy: the moment a pair finishes, its slot takes the next pair of the
// longest-first queue (src/subgradient_solver.hpp:285-324, SIMD twin src/subgradient_solver_simd.hpp:350-395).  Here
// a worker holds `capacity` pair slots.  Every iteration step is one DP launch over the slot table plus one post launch;
// the post kernel appends every pair that finishes to a list, the host reads that list every few steps (without
// draining the stream: the next steps are already queued), and as soon as a tenth of the slots is free it takes that many
// pairs from the device's queue (a COHORT), runs their set-up on a side stream while the iteration steps go on, and
// splices them into the freed slots between two steps.  Every pair carries its own iteration counter
// (PairState::remaining), so pairs of different cohorts iterate side by side.
//
// Slots are positions in BAND GROUPS: the 32 / G pairs that share a warp of the band kernel and its interleaved
// score / trace arrays (lara_gpu.cuh).  A new pair goes into a free slot of a group of its band width whose arrays
// are long enough; pairs that find none open new groups.  Per-pair arrays of a cohort are given back when its last
// pair has finished; the group arrays live as long as the set.
// ---------------------------------------------------------------------------------------------------
struct BandGroup
{
    int G = 0, Q = 0, W = 0, ppw = 0, rows = 0, cols = 0;
    int64_t ov = 0, ot = 0, ou = 0, ol = 0;     // the group's arrays (offsets in ints from the arena base)
    std::vector<int32_t> occ;                   // ppw entries: pair (index in the set's pair arrays) or -1
    int nOcc = 0;
};

struct Cohort
{
    int base = 0, n = 0, live = 0;
    int maxLa = 0, maxLb = 0, maxK = 0;
    std::vector<void *> pooled;                 // per-pair arrays from the stream-ordered pool
    int64_t pooledBytes = 0;
};

struct ResidentSet
{
    lara_ctx & ctx;
    Device & dv;
    Worker & wk;
    cudaStream_t const s;                       // the worker's stream: iteration steps
    cudaStream_t side = nullptr;                // set-up of the next cohort
    cudaEvent_t evSide = nullptr;
    int32_t * arena = nullptr;                  // base of every offset of the set
    PairDesc * dDesc = nullptr;
    PairState * dState = nullptr;
    size_t pairCap = 0;
    std::vector<PairDesc> desc;                 // every pair admitted so far, in admission order
    std::vector<int32_t> cohortOf, groupOf, slotOf;
    std::vector<uint8_t> finishedHost;
    std::vector<Cohort> cohorts;
    std::vector<BandGroup> groups;
    std::map<int, std::vector<int32_t>> fullPairs;      // rows per lane R -> pairs of the full-matrix kernel
    std::vector<void *> keep;                   // allocations that live as long as the set
    int live = 0;
    int64_t liveLanes16 = 0;                    // lanes of the live pairs' fill with blocks of 16 diagonals
    bool have24 = false;
    bool slabFree = true;                       // no live pair has arrays in the worker's slab: the next cohort may take it
    int64_t bytesHeld = 0, keepBytes = 0;       // bytes of the stream-ordered pool the set holds now (all / the groups' arrays)
    // slot table on the device and its launch geometry
    int32_t * dTable = nullptr;
    size_t tableCap = 0;
    size_t nBigClasses = 0;                     // band classes with a launch of their own: bandLaunches[0 .. nBigClasses)
    int nSlots = 0, nPost = 0;                  // slots of the DP classes; pairs of the post kernel (listed after the slots)
    std::vector<BandLaunch> bandLaunches;
    std::vector<std::tuple<int, int, int>> fullLaunches;        // R, offset, count
    int maxLa = 1, maxLb = 1, smemCand = 32;
    // finished list: [0] count, [1 ..] pairs
    int32_t * dFinished = nullptr;

    ResidentSet(lara_ctx & c, Device & d, Worker & w) : ctx(c), dv(d), wk(w), s(w.stream) {}

    void init(size_t maxPairs)
    {
        pairCap = std::max<size_t>(maxPairs, 1);
        CK(malloc_async_retry((void **)&dDesc, sizeof(PairDesc) * pairCap, s));
        CK(malloc_async_retry((void **)&dState, sizeof(PairState) * pairCap, s));
        CK(malloc_async_retry((void **)&dFinished, sizeof(int32_t) * (pairCap + 1), s));
        CK(cudaMemsetAsync(dFinished, 0, sizeof(int32_t), s));
        if (!wk.side)
        {
            CK(cudaStreamCreateWithFlags(&wk.side, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&wk.evSide, cudaEventDisableTiming));
        }
        side = wk.side; evSide = wk.evSide;
        arena = wk.slab ? wk.slab : reinterpret_cast<int32_t *>(dDesc);
    }

    void release()
    {
        // everything the set still holds; the streams are idle (or the device is in an error state) when this runs
        cudaStreamSynchronize(s);
        if (side) cudaStreamSynchronize(side);
        for (Cohort & c : cohorts) for (void * p : c.pooled) cudaFreeAsync(p, s);
        for (void * p : keep) cudaFreeAsync(p, s);
        if (dTable) cudaFreeAsync(dTable, s);
        if (dFinished) cudaFreeAsync(dFinished, s);
        if (dState) cudaFreeAsync(dState, s);
        if (dDesc) cudaFreeAsync(dDesc, s);
        cohorts.clear(); keep.clear();
        dTable = nullptr; dFinished = nullptr; dState = nullptr; dDesc = nullptr; side = nullptr; evSide = nullptr;
        cudaGetLastError();
    }

    // ---- set-up of the pairs [begin, end) of the device's queue on stream st ---------------------------------------
    // Blocks the host twice (edge filter -> E, counts -> D); on return the last set-up kernel is queued on st.  With
    // `pilot` only the stages up to the counts run and the densities are returned.
    void admit(size_t begin, size_t end, cudaStream_t st, double slabHintBytes, Pilot * pilot = nullptr,
               OrientedPair const * pilotPairs = nullptr)
    {
        int const n = (int)(end - begin);
        int const base = (int)desc.size();
        if ((size_t)base + (size_t)n > pairCap) throw std::runtime_error("resident set: pair arrays too small");
        bool const useSlab = slabHintBytes > 0 && slabFree && !pilot;
        if (useSlab && cohorts.empty() && (double)wk.slabInts * 4 < slabHintBytes)
        {
            CK(cudaStreamSynchronize(s));
            CK(cudaFree(wk.slab));
            wk.slab = nullptr; wk.slabInts = 0;
            int64_t const want = ((int64_t)(slabHintBytes * 1.1 / 4) + 1023) & ~(int64_t)1023;
            CK(cudaMalloc((void **)&wk.slab, (size_t)want * 4));
            wk.slabInts = want;
        }
        if (cohorts.empty() && !pilot) arena = wk.slab ? wk.slab : reinterpret_cast<int32_t *>(dDesc);
        if (useSlab) slabFree = false;
        cohorts.emplace_back();
        Cohort * co = &cohorts.back();
        co->base = base; co->n = n;
        // a region: part of the worker's slab (first cohort) or a block of the stream-ordered pool
        auto region_alloc = [&](Region & rg, int64_t sizeInts, int64_t atInts, std::vector<void *> & owner)
        {
            if (useSlab && wk.slab && atInts + sizeInts <= wk.slabInts)
                rg.ptr = wk.slab + atInts;
            else
            {
                int64_t const b = std::max<int64_t>(sizeInts, 1) * 4;
                CK(malloc_async_retry((void **)&rg.ptr, (size_t)b, st));
                owner.push_back(rg.ptr);
                bytesHeld += b;
                if (&owner == &keep) keepBytes += b;
                else if (&owner == &cohorts.back().pooled) cohorts.back().pooledBytes += b;
            }
            rg.size = sizeInts;
            rg.baseDelta = rg.ptr - arena;
        };
        auto al256 = [](int64_t x) { return (x + 63) & ~(int64_t)63; };
        auto tPrev = std::chrono::steady_clock::now();
        auto lap = [&](char const * what)
        {
            if (!ctx.verbose) return;
            auto const now = std::chrono::steady_clock::now();
            std::fprintf(stderr, "[lara_b200] dev %d cohort %zu+%d %-28s %8.2f ms\n", dv.id, begin, n, what,
                         std::chrono::duration<double, std::milli>(now - tPrev).count());
            tPrev = now;
        };
        DevParams const & dp = ctx.dp;
        desc.resize((size_t)base + (size_t)n);
        cohortOf.resize(desc.size(), (int32_t)cohorts.size() - 1);
        groupOf.resize(desc.size(), -1);
        slotOf.resize(desc.size(), -1);
        finishedHost.resize(desc.size(), 0);
        PairDesc * const cd = desc.data() + base;             // this cohort's descriptors
        PairDesc * const dCd = dDesc + base;
        PairState * const dCs = dState + base;
        // ---- stage A layout: everything that only depends on the lengths -----------------------------
        auto layoutA = [&](bool commit, Region & rg)
        {
            for (int i = 0; i < n; ++i)
            {
                OrientedPair const & op = pilot ? pilotPairs[i] : dv.pairs[begin + (size_t)i];
                PairDesc & d = cd[i];
                if (!commit)
                {
                    std::memset(&d, 0, sizeof(d));
                    d.la = op.la; d.lb = op.lb; d.seqFirst = op.first; d.seqSecond = op.second;
                    d.outIdx = (int32_t)(begin + (size_t)i);
                    d.R = (op.lb + 31) / 32;
                    d.W = d.R <= 8 ? 1 : (d.R <= 16 ? 2 : 4);
                    d.nSteps = (op.la + 31 + 3) & ~3;
                    d.bitWords = (op.lb + 31) / 32;
                    d.kCap = std::min(ctx.nbp[op.first], ctx.nbp[op.second]) + 1;
                    co->maxLa = std::max(co->maxLa, d.la); co->maxLb = std::max(co->maxLb, d.lb); co->maxK = std::max(co->maxK, d.kCap);
                }
                d.offBits = rg.take((int64_t)d.la * d.bitWords);
                d.offRowStart = rg.take(d.la + 1);
                d.offLines = rg.take(d.lb);
                d.offBest = rg.take(d.lb);
                {
                    int64_t capPow2 = 32;                   // the bitonic sort pads to a power of two
                    while (capPow2 < d.kCap) capPow2 <<= 1;
                    d.offCand = rg.take(2 * capPow2);
                }
            }
        };
        int64_t sizeA = 0;
        {
            Region probe;
            layoutA(false, probe);
            sizeA = probe.used;
        }
        lap("layout A (host)");
        Region rA;
        region_alloc(rA, sizeA, 0, co->pooled);
        layoutA(true, rA);
        // edge filter scratch (forward prefix scores): lives only until the filter has run
        Region rS;
        std::vector<void *> scratchOwner;
        int64_t scratchBytes = 0;
        {
            int64_t sizeS = 0;
            for (int i = 0; i < n; ++i) sizeS += align4((int64_t)cd[i].la * cd[i].lb + 6ll * (cd[i].la + 1));
            region_alloc(rS, sizeS, al256(sizeA), scratchOwner);
            if (!scratchOwner.empty()) scratchBytes = std::max<int64_t>(sizeS, 1) * 4;
            for (int i = 0; i < n; ++i)
                cd[i].offScratch = rS.take((int64_t)cd[i].la * cd[i].lb + 6ll * (cd[i].la + 1));
        }
        CK(cudaMemsetAsync(dCs, 0, sizeof(PairState) * (size_t)n, st));
        CK(cudaMemcpyAsync(dCd, cd, sizeof(PairDesc) * (size_t)n, cudaMemcpyHostToDevice, st));
        lap("alloc A + scratch");
        cudaEvent_t evS0 = wk.events[0], evS1 = wk.events[1];
        bool const timed = cohorts.size() == 1 && !pilot;       // the first cohort's set-up is timed with events
        if (timed) CK(cudaEventRecord(evS0, st));
        // ---- stage 1: edge filter -> E ------------------------------------------------------------------
        CK(launch_edge_filter(n, dCd, dCs, arena, dv.in, dp, co->maxLb, st));
        CK(cudaMemcpyAsync(cd, dCd, sizeof(PairDesc) * (size_t)n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        for (void * f : scratchOwner) CK(cudaFreeAsync(f, st));
        bytesHeld -= scratchBytes;
        lap("edge filter (sync)");
        wk.stats.kernel_launches += 1;
        // ---- band geometry of every pair; slots in existing groups, new groups for the rest ---------------------
        bool allow24 = false;
        int64_t batchWarps16 = 0;                               // warps of the set's fill with blocks of 16 diagonals
        {
            int64_t fit = 0, lanes16 = liveLanes16;
            for (int i = 0; i < n; ++i)
            {
                choose_band(ctx, cd[i], true);
                fit += cd[i].bandW == 24;
                lanes16 += std::max(1, cd[i].bandW / 32);
            }
            batchWarps16 = (lanes16 + 31) / 32;
            // the 24-diagonal shape exists as one thread per pair only: used when the set runs in the throughput form and
            // enough pairs fit it to make a class worth its while
            allow24 = ctx.band24 && batchWarps16 >= ctx.bandMinWarps && ctx.bandForceQ == 0 &&
                      (have24 || (fit + 31) / 32 >= ctx.band24MinWarps);
        }
        std::map<int, std::vector<int32_t>> byW;               // band width -> pairs (cohort-local index) still without a slot
        for (int i = 0; i < n; ++i)
        {
            choose_band(ctx, cd[i], allow24);
            if (cd[i].bandW > 0) byW[cd[i].bandW].push_back(i);
        }
        if (ctx.verbose)
        {
            int hist[9] = {0}, fine[4] = {0};
            for (int i = 0; i < n; ++i)
            {
                PairDesc const & d = cd[i];
                int const need = (d.E > 0 && d.dMin <= d.dMax) ? std::max(d.dMax + ctx.bandMargin, 0) - std::min(d.dMin - ctx.bandMargin, 0) + 1 : 1;
                int b = 0;
                while (b < 8 && (8 << b) < need) ++b;
                ++hist[b];
                if (need > 16 && need <= 32) ++fine[(need - 17) / 4];
            }
            std::fprintf(stderr, "[lara_b200]   band need histogram (<=8,16,32,64,128,256,512,1024,more):");
            for (int b = 0; b < 9; ++b) std::fprintf(stderr, " %d", hist[b]);
            std::fprintf(stderr, "; 17-20 21-24 25-28 29-32: %d %d %d %d\n", fine[0], fine[1], fine[2], fine[3]);
        }
        size_t const firstNewGroup = groups.size();
        int reused = 0;
        if (!pilot && firstNewGroup > 0)
        {
            // free slots of the groups that are there, first fit (groups were opened longest first)
            std::map<int, std::vector<int>> vacant;                // width -> groups with a free slot
            for (size_t g = 0; g < groups.size(); ++g)
                if (groups[g].nOcc < groups[g].ppw) vacant[groups[g].W].push_back((int)g);
            for (auto & kv : byW)
            {
                auto it = vacant.find(kv.first);
                if (it == vacant.end()) continue;
                std::vector<int32_t> rest;
                for (int32_t i : kv.second)
                {
                    PairDesc & d = cd[i];
                    bool placed = false;
                    for (int g : it->second)
                    {
                        BandGroup & bg = groups[(size_t)g];
                        if (bg.nOcc >= bg.ppw || bg.rows < d.lb + 1 || bg.cols < d.la + 1) continue;
                        int j = 0;
                        while (bg.occ[(size_t)j] >= 0) ++j;
                        bg.occ[(size_t)j] = base + i; ++bg.nOcc;
                        d.bandG = bg.G; d.bandQ = bg.Q; d.bandSlot = j;
                        d.offBVals = bg.ov; d.offBTrace = bg.ot; d.offU = bg.ou; d.offL = bg.ol;
                        groupOf[(size_t)(base + i)] = g; slotOf[(size_t)(base + i)] = j;
                        placed = true; ++reused;
                        break;
                    }
                    if (!placed) rest.push_back(i);
                }
                kv.second.swap(rest);
            }
        }
        for (auto & kv : byW)
        {
            int const W = kv.first;
            std::vector<int32_t> & members = kv.second;
            if (members.empty()) continue;
            std::stable_sort(members.begin(), members.end(), [&](int32_t x, int32_t y) { return cd[x].lb > cd[y].lb; });
            int G, Q;
            choose_band_shape(ctx, W, batchWarps16, ((int64_t)members.size() * std::max(1, W / 32) + 31) / 32, G, Q);
            int const ppw = 32 / G;
            for (size_t k = 0; k < members.size(); k += (size_t)ppw)
            {
                int const cnt = (int)std::min<size_t>((size_t)ppw, members.size() - k);
                BandGroup bg;
                bg.G = G; bg.Q = Q; bg.W = W; bg.ppw = ppw; bg.rows = cd[members[k]].lb + 1;
                bg.occ.assign((size_t)ppw, -1);
                for (int j = 0; j < cnt; ++j)
                {
                    PairDesc & d = cd[members[k + (size_t)j]];
                    bg.cols = std::max(bg.cols, d.la + 1);
                    d.bandG = G; d.bandQ = Q; d.bandSlot = j;
                    bg.occ[(size_t)j] = base + members[k + (size_t)j];
                    groupOf[(size_t)(base + members[k + (size_t)j])] = (int32_t)groups.size();
                    slotOf[(size_t)(base + members[k + (size_t)j])] = j;
                }
                bg.nOcc = cnt;
                groups.push_back(bg);
            }
            if (W == 24) have24 = true;
        }
        // ---- stage B layout: per-edge arrays (with the cohort), band arrays of the new groups (with the set) ---------
        int64_t sizeB = 0, sizeG = 0;
        for (int i = 0; i < n; ++i)
        {
            PairDesc & d = cd[i];
            sizeB += 8 * align4(std::max(d.E, 1)) + align4(d.E + 1);
            if (d.bandW > 0 && d.bandW <= 32) sizeB += align4(4ll * d.la);
            if (d.bandW == 0)
            {
                // no usable band: the single-pass full-matrix kernel up to 640 nt, the row-by-row kernel beyond
                d.dense = (d.R > kMaxR || ctx.forceRows || ctx.rowsMode == 1) ? 1 : 0;
                if (d.dense) sizeB += align4(4ll * (d.la + 1)) + align4(((int64_t)(d.lb + 1) * (d.la + 1) + 3) / 4);
                else sizeB += align4(32ll * ((d.nSteps + 31) / 32)) + 32 + align4(32ll * d.nSteps * d.W);
            }
        }
        for (size_t g = firstNewGroup; g < groups.size(); ++g)
        {
            BandGroup const & bg = groups[g];
            sizeG += align4((int64_t)bg.rows * bg.W * bg.ppw) + 8 +
                     (bg.ppw == 32 ? (int64_t)((bg.rows + 1) & ~1) * 128 + 8 : align4((int64_t)bg.rows * (bg.W / 8) * bg.ppw)) +
                     4ll * bg.ppw * (bg.rows + bg.cols);
        }
        lap("layout B (host)");
        co = &cohorts.back();
        Region rB, rG;
        region_alloc(rB, sizeB, al256(sizeA), co->pooled);
        region_alloc(rG, sizeG, al256(sizeA) + al256(sizeB), pilot ? co->pooled : keep);
        lap("alloc B");
        for (int i = 0; i < n; ++i)
        {
            PairDesc & d = cd[i];
            d.offCell = rB.take(d.E); d.offSeqScore = rB.take(d.E); d.offPv = rB.take(d.E); d.offPerm = rB.take(d.E);
            d.offIoff = rB.take(d.E + 1);
            d.offEdge = rB.take(4ll * d.E);
            if (d.bandW > 0 && d.bandW <= 32) d.offRowRec = rB.take(4ll * d.la);
            if (d.bandW > 0)
            {
                if (!pilot)
                {
                    ++wk.stats.reserved[1];                     // pairs solved by the band kernels
                    if (d.bandQ == 16) ++wk.stats.reserved[2];  // ... of those in blocks of 16 diagonals
                    wk.stats.reserved[3] += d.bandW;            // sum of their band widths
                }
            }
            else if (d.dense)
            {
                d.offFlags = rB.take(4ll * (d.la + 1));                                     // S, V of two rows
                d.offTrace = rB.take(((int64_t)(d.lb + 1) * (d.la + 1) + 3) / 4);           // one decision byte per cell
            }
            else
            {
                d.offFlags = rB.take(32ll * ((d.nSteps + 31) / 32));
                d.offLanePtr = rB.take(32);
                d.offTrace = rB.take(32ll * d.nSteps * d.W);
            }
        }
        for (size_t g = firstNewGroup; g < groups.size(); ++g)
        {
            BandGroup & bg = groups[g];
            bg.ov = rG.take((int64_t)bg.rows * bg.W * bg.ppw);
            // trace: W / 8 ints per row and pair; one thread per pair (ppw 32, W 32 or 24): 16-byte row records, two
            // rows per 32-byte sector
            if (bg.ppw == 32) rG.used = (rG.used + 7) & ~(int64_t)7;
            bg.ot = bg.ppw == 32 ? rG.take((int64_t)((bg.rows + 1) & ~1) * 4 * 32)
                                 : rG.take((int64_t)bg.rows * (bg.W / 8) * bg.ppw);
            bg.ou = rG.take(4ll * bg.ppw * bg.rows);
            bg.ol = rG.take(4ll * bg.ppw * bg.cols);
            for (int j = 0; j < bg.ppw; ++j)
                if (bg.occ[(size_t)j] >= 0)
                {
                    PairDesc & d = desc[(size_t)bg.occ[(size_t)j]];
                    d.offBVals = bg.ov; d.offBTrace = bg.ot; d.offU = bg.ou; d.offL = bg.ol;
                }
        }
        CK(cudaMemcpyAsync(dCd, cd, sizeof(PairDesc) * (size_t)n, cudaMemcpyHostToDevice, st));
        // ---- stage 2: counts -> D, nVals -----------------------------------------------------------------------
        CK(launch_setup_count(n, dCd, dCs, arena, dv.in, dp, st));
        CK(cudaMemcpyAsync(cd, dCd, sizeof(PairDesc) * (size_t)n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        lap("set-up count (sync)");
        wk.stats.kernel_launches += 1;
        if (pilot)
        {
            for (int i = 0; i < n; ++i)
            {
                PairDesc const & d = cd[i];
                pilot->cells += (double)d.la * d.lb;
                pilot->edges += d.E;
                pilot->inters += d.D;
                if (d.bandW > 0)
                {
                    pilot->band = true;
                    int const lo = std::min(d.dMin - ctx.bandMargin, 0), hi = std::max(d.dMax + ctx.bandMargin, 0);
                    if (d.E > 0 && d.dMin <= d.dMax) pilot->bandSlack = std::max(pilot->bandSlack, hi - lo + 1 - (d.la - d.lb + 1));
                }
                // scores + trace + closed-form records of the band, or stream + trace of the full-matrix kernel
                pilot->dpBytes += d.bandW > 0 ? 4.6 * d.bandW * (d.lb + 1) + 16.0 * (d.la + d.lb + 2)
                                 : (d.dense ? 5.0 * (d.la + 1.0) * (d.lb + 1.0) : 4.0 * d.nVals + 128.0 * d.nSteps * d.W);
            }
            return;
        }
        int64_t sizeC = 0;
        for (int i = 0; i < n; ++i) sizeC += 14 * align4(std::max(cd[i].D, 1)) + 20 + align4(std::max(cd[i].nVals, 1));
        Region rC;
        region_alloc(rC, sizeC, al256(sizeA) + al256(sizeB) + al256(sizeG), co->pooled);
        lap("alloc C");
        for (int i = 0; i < n; ++i)
        {
            PairDesc & d = cd[i];
            d.offPart = rC.take(d.D); d.offPpos = rC.take(d.D); d.offSc = rC.take(d.D); d.offDual = rC.take(d.D);
            d.offRev = rC.take(d.D); d.offList = rC.take(d.D); d.offVals = rC.take(d.nVals);
            rC.used = (rC.used + 15) & ~(int64_t)15;          // hot records: rows start on 64-byte fetch units where they can
            d.offInter = rC.take(4ll * d.D);
            d.offInterCold = rC.take(4ll * d.D);
            wk.stats.n_active_edges += d.E; wk.stats.n_interactions += d.D;
            wk.stats.filter_cells += 2ll * d.la * d.lb;
            if (d.bandW == 0) fullPairs[d.dense ? 0 : d.R].push_back(base + i);   // key 0: row-by-row kernel
            liveLanes16 += std::max(1, d.bandW / 32);
        }
        CK(cudaMemcpyAsync(dCd, cd, sizeof(PairDesc) * (size_t)n, cudaMemcpyHostToDevice, st));
        // ---- stage 3: fill ---------------------------------------------------------------------------------------
        CK(launch_setup_fill(n, dCd, dCs, arena, dv.in, dp, st));
        wk.stats.kernel_launches += 1;
        if (timed) CK(cudaEventRecord(evS1, st));
        co->live = n;
        live += n;
        wk.stats.n_pairs += n;
        if (cohorts.size() == 1)
        {
            // algorithmic HBM bytes of one DP launch with every pair of the first cohort live (bench.py: traffic probe)
            int64_t bytes = 0, pairs = 0;
            for (int i = 0; i < n; ++i)
                if (cd[i].bandW > 0) { bytes += (int64_t)(4.5 * cd[i].bandW * cd[i].lb + 16.0 * (cd[i].la + cd[i].lb + 2)); ++pairs; }
            wk.stats.reserved[10] = bytes; wk.stats.reserved[11] = pairs;
        }
        else
            wk.stats.reserved[12] += n;                       // pairs taken into a running set
        if (ctx.verbose)
            std::fprintf(stderr, "[lara_b200] dev %d cohort %zu+%d: %d into free slots, %zu new groups, %d pairs live\n", dv.id, begin, n,
                         reused, groups.size() - firstNewGroup, live);
    }

    // Every pair of the set has finished: forget its groups, so that the next cohort starts from the worker's slab like
    // the first one did (no pair's arrays are in use any more; st: the stream the next set-up will run on).
    void begin_new_set(cudaStream_t st)
    {
        for (Cohort & c : cohorts)
        {
            for (void * f : c.pooled) CK(cudaFreeAsync(f, s));
            c.pooled.clear();
            bytesHeld -= c.pooledBytes; c.pooledBytes = 0;
        }
        for (void * f : keep) CK(cudaFreeAsync(f, s));
        keep.clear();
        bytesHeld -= keepBytes; keepBytes = 0;
        groups.clear();
        for (int32_t & g : groupOf) g = -1;
        fullPairs.clear();
        have24 = false;
        liveLanes16 = 0;
        slabFree = true;
        // the set-up kernels of the next cohort write into the slab: after everything queued on the worker's stream
        CK(cudaEventRecord(evSide, s));
        CK(cudaStreamWaitEvent(st, evSide, 0));
    }

    // ---- slot table: the occupied groups class by class (most slots first), then the full-matrix classes -----------
    void rebuild_table()
    {
        std::map<std::pair<int, int>, std::vector<int>> byShape;        // (G, Q) -> groups with a pair in them
        for (size_t g = 0; g < groups.size(); ++g)
            if (groups[g].nOcc > 0) byShape[{groups[g].G, groups[g].Q}].push_back((int)g);
        std::vector<std::tuple<int64_t, int, int>> order;              // slots, G, Q
        for (auto const & kv : byShape) order.emplace_back((int64_t)kv.second.size() * (32 / kv.first.first), kv.first.first, kv.first.second);
        std::sort(order.begin(), order.end(), [](auto const & x, auto const & y) { return std::get<0>(x) > std::get<0>(y); });
        std::vector<int32_t> table;
        bandLaunches.clear(); fullLaunches.clear();
        for (auto const & o : order)
        {
            int const G = std::get<1>(o), Q = std::get<2>(o);
            int const off = (int)table.size();
            for (int g : byShape[{G, Q}])
                table.insert(table.end(), groups[(size_t)g].occ.begin(), groups[(size_t)g].occ.end());
            bandLaunches.push_back(BandLaunch{G, Q, (int)table.size() - off, off});
        }
        {
            // classes (other than the largest) that fill a good part of the machine by themselves come first: they get
            // launches of their own (enqueue_step)
            nBigClasses = 0;
            for (size_t k = 1; k < bandLaunches.size(); ++k)
                if ((int64_t)bandLaunches[k].nSlots * bandLaunches[k].G / 32 >= ctx.bandSplitWarps)
                    std::swap(bandLaunches[nBigClasses++], bandLaunches[k]);
            // (the swap moved the largest class behind the big ones when there are any; order inside the rest does not matter)
        }
        for (auto & kv : fullPairs)
        {
            std::vector<int32_t> & v = kv.second;
            v.erase(std::remove_if(v.begin(), v.end(), [&](int32_t p) { return finishedHost[(size_t)p] != 0; }), v.end());
            if (v.empty()) continue;
            fullLaunches.emplace_back(kv.first, (int)table.size(), (int)v.size());
            table.insert(table.end(), v.begin(), v.end());
        }
        nSlots = (int)table.size();
        // the post kernel takes the pairs in admission order (== the order of their arrays in memory), one warp each
        {
            std::vector<int32_t> inOrder;
            inOrder.reserve(table.size());
            for (int32_t p : table) if (p >= 0) inOrder.push_back(p);
            std::sort(inOrder.begin(), inOrder.end());
            nPost = (int)inOrder.size();
            table.insert(table.end(), inOrder.begin(), inOrder.end());
        }
        if (table.size() > tableCap)
        {
            if (dTable) CK(cudaFreeAsync(dTable, s));
            tableCap = table.size() + table.size() / 4 + 1024;
            CK(malloc_async_retry((void **)&dTable, sizeof(int32_t) * tableCap, s));
        }
        if (!table.empty())
        {
            // through a page-locked buffer: a copy from pageable memory would drain the stream first
            if (wk.hTableCap < table.size())
            {
                if (wk.hTable) { CK(cudaStreamSynchronize(s)); CK(cudaFreeHost(wk.hTable)); }
                wk.hTableCap = table.size() + table.size() / 4 + 1024;
                CK(cudaMallocHost((void **)&wk.hTable, sizeof(int32_t) * wk.hTableCap));
            }
            if (!wk.evTable) CK(cudaEventCreateWithFlags(&wk.evTable, cudaEventDisableTiming));
            else CK(cudaEventSynchronize(wk.evTable));                  // the previous upload has left the buffer
            std::memcpy(wk.hTable, table.data(), sizeof(int32_t) * table.size());
            CK(cudaMemcpyAsync(dTable, wk.hTable, sizeof(int32_t) * table.size(), cudaMemcpyHostToDevice, s));
            CK(cudaEventRecord(wk.evTable, s));
        }
        maxLa = 1; maxLb = 1;
        int maxK = 1;
        for (Cohort const & c : cohorts)
            if (c.live > 0) { maxLa = std::max(maxLa, c.maxLa); maxLb = std::max(maxLb, c.maxLb); maxK = std::max(maxK, c.maxK); }
        // shared memory candidates of the matching: as many as fit next to the line arrays (cap 2048)
        smemCand = 32;
        while (smemCand < maxK && smemCand < ctx.postSmemCand) smemCand <<= 1;
        if (ctx.verbose)
        {
            for (BandLaunch const & b : bandLaunches)
                std::fprintf(stderr, "[lara_b200]   DP class band G=%d Q=%d (W=%d): %d slots\n", b.G, b.Q, 2 * b.G * b.Q, b.nSlots);
            for (auto const & f : fullLaunches)
                std::fprintf(stderr, "[lara_b200]   DP class full matrix %s=%d: %d pairs\n", std::get<0>(f) ? "R" : "row by row, R", std::get<0>(f), std::get<2>(f));
        }
    }

    // a pair the device reported as finished: its slot and, with the last pair of its cohort, the cohort's arrays are free
    void on_finished(int p)
    {
        if (p < 0 || (size_t)p >= desc.size() || finishedHost[(size_t)p]) return;
        finishedHost[(size_t)p] = 1;
        --live;
        liveLanes16 -= std::max(1, desc[(size_t)p].bandW / 32);
        int const g = groupOf[(size_t)p];
        if (g >= 0)
        {
            BandGroup & bg = groups[(size_t)g];
            bg.occ[(size_t)slotOf[(size_t)p]] = -1;
            --bg.nOcc;
        }
        Cohort & c = cohorts[(size_t)cohortOf[(size_t)p]];
        if (--c.live == 0)
        {
            for (void * f : c.pooled) CK(cudaFreeAsync(f, s));
            c.pooled.clear();
            bytesHeld -= c.pooledBytes; c.pooledBytes = 0;
        }
    }

    // one iteration step: DP launch(es) + post launch
    void enqueue_step(cudaEvent_t e0, cudaEvent_t e1, cudaEvent_t e2)
    {
        DevParams const & dp = ctx.dp;
        CK(cudaEventRecord(e0, s));
        // full-matrix classes (pairs without a usable band) on side streams, all band classes in one launch
        // (LARA_B200_SPLIT=1: classes that fill a good part of the machine on their own run as launches of their own beside
        // it; measured on C3 within 1-2 % of the single launch, either way, from box to box: off)
        size_t const nSplit = ctx.splitLaunch ? nBigClasses : 0;
        size_t const nSide = fullLaunches.size() + nSplit;
        if (nSide)
        {
            if (!wk.evFork) CK(cudaEventCreateWithFlags(&wk.evFork, cudaEventDisableTiming));
            while (wk.aux.size() < nSide)
            {
                cudaStream_t st; cudaEvent_t ev;
                CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
                wk.aux.push_back(st); wk.auxJoin.push_back(ev);
            }
            CK(cudaEventRecord(wk.evFork, s));
            for (size_t k = 0; k < fullLaunches.size(); ++k)
            {
                CK(cudaStreamWaitEvent(wk.aux[k], wk.evFork, 0));
                if (std::get<0>(fullLaunches[k]) == 0)
                    CK(launch_gotoh_rows(ctx.rowsMode, std::get<2>(fullLaunches[k]), dTable + std::get<1>(fullLaunches[k]), dDesc, dState, arena,
                                         dp.go, dp.ge, wk.aux[k]));
                else
                    CK(launch_gotoh_fill(std::get<0>(fullLaunches[k]), std::get<2>(fullLaunches[k]), dTable + std::get<1>(fullLaunches[k]),
                                         dDesc, dState, arena, dv.negBuf, dp.go, dp.ge, wk.aux[k]));
                CK(cudaEventRecord(wk.auxJoin[k], wk.aux[k]));
                wk.stats.kernel_launches += 1; wk.stats.dp_launches += 1;
            }
            for (size_t k = 0; k < nSplit; ++k)
            {
                size_t const a = fullLaunches.size() + k;
                CK(cudaStreamWaitEvent(wk.aux[a], wk.evFork, 0));
                CK(launch_gotoh_band_multi(1, &bandLaunches[k], dTable, dDesc, dState, arena, dp.go, dp.ge, wk.aux[a]));
                CK(cudaEventRecord(wk.auxJoin[a], wk.aux[a]));
                wk.stats.kernel_launches += 1; wk.stats.dp_launches += 1;
            }
        }
        if (!bandLaunches.empty())
        {
            // bandLaunches: the nBigClasses classes with launches of their own first, then the largest class and the small ones
            CK(launch_gotoh_band_multi((int)(bandLaunches.size() - nSplit), bandLaunches.data() + nSplit, dTable, dDesc, dState, arena,
                                       dp.go, dp.ge, s));
            wk.stats.kernel_launches += 1; wk.stats.dp_launches += 1;
        }
        for (size_t k = 0; k < nSide; ++k) CK(cudaStreamWaitEvent(s, wk.auxJoin[k], 0));
        CK(cudaEventRecord(e1, s));
        if (ctx.postInOrder)
            CK(launch_iterate_post(nPost, dTable + nSlots, dDesc, dState, arena, dp, dv.dOutHdr, dv.dOutLines, dv.dOutOff, dFinished,
                                   maxLa, maxLb, smemCand, dv.in, s));
        else
            CK(launch_iterate_post(nSlots, dTable, dDesc, dState, arena, dp, dv.dOutHdr, dv.dOutLines, dv.dOutOff, dFinished,
                                   maxLa, maxLb, smemCand, dv.in, s));
        wk.stats.kernel_launches += 1;
        CK(cudaEventRecord(e2, s));
        wk.stats.reserved[7] += 1;
        wk.stats.reserved[8] += nSlots;
    }
};

// Pairs per batch from the free device memory and a per-pair estimate (exact for the length dependent
// part, heuristic for the edge/interaction arrays whose sizes are only known after set-up).
// Bytes of device memory a pair holds while it is resident, estimated from its lengths and the densities measured on a
// sample of the device's pairs (run_device) with a safety factor: bitmap C / 8 + per edge 36 + per interaction 56 + DP
// arrays + lines, candidates, row starts.  The edge filter's scratch (4 C) is held during the set-up of a cohort only.
// `scale`: correction of the resident part from the cohorts a worker has admitted so far (run_worker).
// The DP arrays are estimated per pair, not from a density: the band must hold the diagonals 0 .. la - lb (both corners
// of the matrix lie on active diagonals) plus what the sample needed beyond them, rounded up to a band width that exists;
// a mean over all pairs is off by a factor of 8 either way for the longest and the most unequal pairs of C5 (the complete
// C5 ran out of memory on it).
double pair_bytes(Device const & dv, OrientedPair const & op, double scale = 1.0)
{
    double const cells = std::max(dv.pilot.cells, 1.0);
    double const ePerCell = 1.3 * dv.pilot.edges / cells, dPerCell = 1.3 * dv.pilot.inters / cells;
    double const C = (double)op.la * op.lb;
    double dp = 1.3 * dv.pilot.dpBytes / cells * C;
    if (dv.pilot.band)
    {
        int const need = op.la - op.lb + 1 + dv.pilot.bandSlack;
        int W = 32;
        while (W < need) W <<= 1;
        bool const full = W > 1024 || (op.lb <= 32 * kMaxR && (double)W * (op.lb + 1) >= C);
        // band: scores + trace + closed-form records, 15 % for the unequal rows and the empty slots of a group;
        // full matrix: the denser of the two full-matrix forms
        dp = full ? 5.0 * (op.la + 1.0) * (op.lb + 1.0)
                  : 1.15 * (4.6 * W * (op.lb + 1.0) + 16.0 * (op.la + op.lb + 2.0));
    }
    return 4.0 * C + scale * (C / 8 + (36.0 * ePerCell + 56.0 * dPerCell) * C + dp + 64.0 * (op.la + op.lb) + 4096);
}

// What one of the device's nWorkers workers may hold: its share of 55 % of the device (the stream-ordered pool keeps
// freed blocks per stream, so the footprint runs ahead of the bytes held; at 70 % the complete C5 met "out of memory"), but not more than what is free
// now plus the slabs the workers already hold (other contexts or processes may live on the same device).
double worker_budget(lara_ctx const & ctx, Device & dv, int nWorkers)
{
    if (ctx.memBudget > 0) return ctx.memBudget / (double)nWorkers;
    size_t freeB = 0, totalB = 0;
    cudaMemGetInfo(&freeB, &totalB);
    double held = 0;
    for (Worker const & w : dv.workers) held += 4.0 * (double)w.slabInts;
    return std::min(0.55 * (double)totalB, 0.9 * ((double)freeB + held)) / (double)nWorkers;
}

size_t batch_limit(lara_ctx & ctx, Device & dv, size_t begin, int nWorkers, double & bytes, double budgetScale = 1.0,
                   size_t * slotLimit = nullptr)
{
    double const budget = budgetScale * worker_budget(ctx, dv, nWorkers);
    double used = 0;
    size_t k = begin;
    size_t hardCap = ctx.batchPairsCap ? ctx.batchPairsCap : ctx.batchPairsDefault;
    if (!ctx.batchPairsCap)
    {
        // equal shares, a multiple of the worker count of them, so that no worker ends with a small set alone
        size_t const N = dv.pairs.size();
        size_t nB = (N + hardCap - 1) / hardCap;
        nB = (nB + (size_t)nWorkers - 1) / (size_t)nWorkers * (size_t)nWorkers;
        hardCap = std::max<size_t>((N + nB - 1) / nB, std::min<size_t>(N, 1024));
    }
    while (k < dv.pairs.size() && k - begin < hardCap)
    {
        double const per = pair_bytes(dv, dv.pairs[k]);
        if (used + per > budget && k > begin) break;
        used += per;
        ++k;
    }
    bytes = used;
    if (slotLimit) *slotLimit = hardCap;
    return k;
}

// One worker: a resident set that is refilled from the device's queue until the queue is empty and the last pair has
// finished.
void run_worker(lara_ctx & ctx, Device & dv, Worker & wk, int nWorkers)
{
    ResidentSet rs(ctx, dv, wk);
    try
    {
        CK(cudaSetDevice(dv.id));
        cudaStream_t const s = wk.stream;
        size_t begin, end;
        double bytes = 0;
        double const budget = worker_budget(ctx, dv, nWorkers);
        size_t slotLimit = 0, roundPairs = 0;       // most pairs live at once; pairs of a set that takes the whole budget
        double estScale = 1.0;                      // resident bytes per pair: measured / estimated, over the cohorts so far
        {
            std::lock_guard<std::mutex> lock(dv.queueMutex);
            begin = dv.nextBegin;
            end = begin < dv.pairs.size() ? batch_limit(ctx, dv, begin, nWorkers, bytes, 1.0, &slotLimit) : begin;
            roundPairs = end - begin;
            if (end > begin && ctx.refillMode != 0 &&
                (ctx.refillMode == 2 || dv.pairs.size() - begin > 3 * (end - begin) * (size_t)nWorkers))
            {
                // The queue is long: slots will be refilled as they free up.  A cohort gives its arrays back when its
                // last pair has finished, so the set holds somewhat more than its live pairs' arrays: the first cohort
                // leaves room for that, and its arrays come from the pool (arrays in the worker's slab could only be
                // given back when the whole set has drained).
                end = batch_limit(ctx, dv, begin, nWorkers, bytes, ctx.refillFirstShare);
                bytes = 0;
            }
            dv.nextBegin = end;
        }
        if (end > begin)
        {
            size_t const capacity = std::max(end - begin, slotLimit);
            rs.init(dv.pairs.size());
            // estimate of the resident bytes of the pairs [b, e) with scale 1 (under the queue's order: no lock needed)
            auto resident_estimate = [&](size_t b, size_t e)
            {
                double x = 0;
                for (size_t k = b; k < e; ++k) x += pair_bytes(dv, dv.pairs[k]) - 4.0 * (double)dv.pairs[k].la * dv.pairs[k].lb;
                return x;
            };
            // after a cohort from the pool: how its bytes compare with the estimate (10 % on top; never below the a-priori one)
            auto calibrate = [&](size_t b, size_t e, int64_t heldBefore)
            {
                if (e - b < 256) return;
                double const ratio = (double)(rs.bytesHeld - heldBefore) / std::max(resident_estimate(b, e), 1.0);
                // one-sided: the queue is sorted by length, and what the longest pairs hold per estimated byte says nothing
                // about the pairs behind them (complete C5: 0.13 for the first 2,000 pairs, above 1 for the next 15,000, and
                // "out of memory" when the first figure sized the second cohort); a set that outgrew its estimate does count
                estScale = std::min(1.5, std::max(1.0, 1.1 * ratio));
                if (ctx.verbose)
                    std::fprintf(stderr, "[lara_b200] dev %d cohort %zu+%zu: %.2f GB resident, %.2f of the estimate; set holds %.2f of %.2f GB\n",
                                 dv.id, b, e - b, (double)(rs.bytesHeld - heldBefore) / 1e9, ratio, (double)rs.bytesHeld / 1e9, budget / 1e9);
            };
            rs.admit(begin, end, s, bytes);
            if (bytes == 0) calibrate(begin, end, 0);
            rs.rebuild_table();
            int const nIter = ctx.P.num_iterations;
            // finished list window read by the host after every block of steps: [0] count, [1 ..] the entries from
            // the host's read position on
            size_t const window = std::min<size_t>(dv.pairs.size(), 65536);
            if (wk.hFinishedCap < 2 * (window + 1))
            {
                if (wk.hFinished) CK(cudaFreeHost(wk.hFinished));
                CK(cudaMallocHost((void **)&wk.hFinished, sizeof(int32_t) * 2 * (window + 1)));
                wk.hFinishedCap = 2 * (window + 1);
            }
            if (!wk.evBlock[0]) for (int k = 0; k < 2; ++k) CK(cudaEventCreateWithFlags(&wk.evBlock[k], cudaEventDisableTiming));
            size_t readPos = 0;                             // finished entries the host has processed
            size_t winStart[2] = {0, 0};
            int steps = 0, block = 0, nextBlockLen = 1;
            // every set may run numIterations steps, and its end is noticed up to two blocks of steps late
            int64_t const maxSteps = ((int64_t)nIter + 64) * (int64_t)(20 * dv.pairs.size() / std::max<size_t>(end - begin, 1) + 3) + 64;
            bool more = true;                               // the queue may still hold pairs
            bool pendingSplice = false;
            int outstanding = -1;                           // buffer of the block whose results have not been read yet
            for (;;)
            {
                // ---- queue a block of steps ----------------------------------------------------------------------
                int const len = nextBlockLen;
                nextBlockLen = std::min(16, nextBlockLen * 2);
                while (wk.events.size() < 2 + 3 * (size_t)(steps + len))
                {
                    cudaEvent_t e;
                    CK(cudaEventCreate(&e));
                    wk.events.push_back(e);
                }
                if (pendingSplice)
                {
                    // the pairs whose set-up ran on the side stream join between two steps
                    CK(cudaStreamWaitEvent(s, rs.evSide, 0));
                    rs.rebuild_table();
                    pendingSplice = false;
                }
                for (int k = 0; k < len; ++k, ++steps)
                    rs.enqueue_step(wk.events[2 + 3 * (size_t)steps], wk.events[2 + 3 * (size_t)steps + 1], wk.events[2 + 3 * (size_t)steps + 2]);
                int const cur = block & 1;
                int32_t * const hb = wk.hFinished + (size_t)cur * (window + 1);
                winStart[cur] = readPos;
                CK(cudaMemcpyAsync(hb, rs.dFinished, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
                CK(cudaMemcpyAsync(hb + 1, rs.dFinished + 1 + readPos, sizeof(int32_t) * std::min(window, rs.pairCap - readPos),
                                   cudaMemcpyDeviceToHost, s));
                CK(cudaEventRecord(wk.evBlock[cur], s));
                ++block;
                // ---- results of the block before this one (the device is busy with the block just queued) ---------------
                auto read_block = [&](int b)
                {
                    CK(cudaEventSynchronize(wk.evBlock[b]));
                    int32_t const * const rb = wk.hFinished + (size_t)b * (window + 1);
                    size_t const count = (size_t)rb[0];
                    size_t const from = std::max(readPos, winStart[b]);
                    size_t const to = std::min(count, winStart[b] + window);
                    for (size_t k = from; k < to; ++k) rs.on_finished(rb[1 + (k - winStart[b])]);
                    if (to > readPos) readPos = to;
                };
                if (outstanding >= 0) read_block(outstanding);
                outstanding = cur;
                // ---- refill: a tenth of the slots free (or nothing left to iterate) and pairs waiting -----------------------
                if (more && !pendingSplice)
                {
                    size_t const freeSlots = capacity > (size_t)rs.live ? capacity - (size_t)rs.live : 0;
                    // Eager refill keeps every slot busy, but a pair taken in late runs up to numIterations steps beyond
                    // the pairs that were there before it: when the queue is about to run dry that only stretches the
                    // tail (measured on C3, where three pairs in four use all 500 iterations: 2,110 steps instead of
                    // 2,000).  So slots are refilled as they free up while the queue still holds more than two full
                    // rounds of all workers' sets, and set by set (when the set has drained) after that.
                    size_t queued;
                    {
                        std::lock_guard<std::mutex> lock(dv.queueMutex);
                        queued = dv.pairs.size() - std::min(dv.pairs.size(), dv.nextBegin);
                    }
                    roundPairs = std::max(roundPairs, (size_t)rs.live);
                    bool const eager = ctx.refillMode == 2 || (ctx.refillMode == 1 && queued > 2 * roundPairs * (size_t)nWorkers);
                    size_t const minCohort = std::max<size_t>((size_t)rs.live / 10, 1);
                    if ((eager && freeSlots >= minCohort) || rs.live == 0)
                    {
                        size_t b2, e2;
                        bool wait = false;
                        {
                            // as many of the free slots as the worker's memory budget has room for, counted from what
                            // the set holds now (a drained set gives everything back first)
                            std::lock_guard<std::mutex> lock(dv.queueMutex);
                            b2 = dv.nextBegin;
                            size_t const last = std::min(dv.pairs.size(), b2 + std::max<size_t>(freeSlots, 1));
                            double const held = rs.live == 0 ? 0.0 : (double)rs.bytesHeld;
                            double est = 0;
                            for (e2 = b2; e2 < last; ++e2)
                            {
                                double const per = pair_bytes(dv, dv.pairs[e2], estScale);
                                if (held + est + per > budget && (e2 > b2 || rs.live > 0)) break;
                                est += per;
                            }
                            // too little room for a cohort worth its set-up: wait for running cohorts to give memory back
                            wait = rs.live > 0 && e2 < last && e2 - b2 < minCohort;
                            if (wait) e2 = b2;
                            dv.nextBegin = e2;
                        }
                        if (wait) {}
                        else if (e2 > b2)
                        {
                            // the slots are known to be free: their pairs were reported by a block that has completed
                            double bytes2 = 0;
                            if (rs.live == 0)
                            {
                                // a drained set: the next one is laid out in the slab like the first
                                rs.begin_new_set(rs.side);
                                bytes2 = (double)wk.slabInts * 4;       // what does not fit the slab comes from the pool
                            }
                            int64_t const heldBefore = rs.bytesHeld;
                            rs.admit(b2, e2, rs.side, bytes2);
                            if (bytes2 == 0) calibrate(b2, e2, heldBefore);
                            CK(cudaEventRecord(rs.evSide, rs.side));
                            pendingSplice = true;
                        }
                        else
                            more = false;
                    }
                }
                if (rs.live == 0 && !pendingSplice && !more)
                {
                    // every pair is accounted for; the block just queued finds nothing to do
                    if (outstanding >= 0) read_block(outstanding);
                    break;
                }
                if (steps > maxSteps) throw std::runtime_error("pairs did not finish within the iteration limit");
            }
            CK(cudaStreamSynchronize(s));
            // ---- per-pair outcome: error flags, iteration counts, work counters --------------------------------------------
            size_t const nAll = rs.desc.size();
            std::vector<PairState> hState(nAll);
            CK(cudaMemcpyAsync(hState.data(), rs.dState, sizeof(PairState) * nAll, cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            for (size_t i = 0; i < nAll; ++i)
            {
                if (hState[i].error)
                    throw std::runtime_error("internal invariant violated on the device (code " + std::to_string(hState[i].error) +
                                             ") in pair (" + std::to_string(rs.desc[i].seqFirst) + ", " + std::to_string(rs.desc[i].seqSecond) + ")");
                if (!hState[i].done)
                    throw std::runtime_error("pair did not finish within numIterations");
                int64_t const its = ctx.P.num_iterations - hState[i].remaining;
                PairDesc const & d = rs.desc[i];
                wk.stats.sum_iterations += its;
                wk.stats.reserved[9] += its;
                wk.stats.dp_cells += its * (int64_t)d.la * d.lb;
                // algorithmic HBM bytes of the DP launches: band: 4 B score + 0.5 B trace per band cell and
                // 16 B closed-form record per row and column; full matrix: 1.3 B per cell (sparse stream + trace)
                wk.stats.reserved[4] += d.bandW > 0 ? its * (int64_t)(4.5 * d.bandW * d.lb + 16.0 * (d.la + d.lb + 2))
                                                    : its * (int64_t)(1.3 * d.la * d.lb);
                wk.stats.reserved[5] += d.bandW > 0 ? its * (int64_t)d.bandW * d.lb : its * (int64_t)d.la * d.lb;   // cells computed
            }
            wk.stats.ms_setup += ev_ms(wk.events[0], wk.events[1]);
            for (int it = 0; it < steps; ++it)
            {
                wk.stats.ms_dp += ev_ms(wk.events[2 + 3 * (size_t)it], wk.events[2 + 3 * (size_t)it + 1]);
                wk.stats.ms_bounds += ev_ms(wk.events[2 + 3 * (size_t)it + 1], wk.events[2 + 3 * (size_t)it + 2]);
            }
            wk.stats.reserved[6] += (int64_t)rs.cohorts.size();           // cohorts (the first one + refills)
        }
        rs.release();
        CK(cudaEventRecord(wk.evDone, wk.stream));
        CK(cudaStreamSynchronize(wk.stream));
    }
    catch (std::exception const & e)
    {
        wk.err = e.what();
        cudaGetLastError();
        rs.release();
    }
}

void run_device(lara_ctx & ctx, Device & dv)
{
    try
    {
        CK(cudaSetDevice(dv.id));
        std::memset(&dv.stats, 0, sizeof(dv.stats));
        dv.err.clear();
        // few pairs: one worker is enough (and keeps the small-case latency low)
        int nW = (int)std::min<size_t>((size_t)ctx.nWorkers, (dv.pairs.size() + ctx.workerMinPairs - 1) / ctx.workerMinPairs);
        nW = std::max(1, std::min(nW, (int)dv.workers.size()));
        dv.nextBegin = 0;
        for (Worker & wk : dv.workers) { std::memset(&wk.stats, 0, sizeof(wk.stats)); wk.err.clear(); }
        // pilot: set-up stages on a sample of the pairs -> densities for the batch sizes
        if (dv.pilot.cells == 0 && !dv.pairs.empty())
        {
            size_t const nS = std::min<size_t>(dv.pairs.size(), 256);
            std::vector<OrientedPair> sample(nS);
            for (size_t i = 0; i < nS; ++i) sample[i] = dv.pairs[i * dv.pairs.size() / nS];
            lara_stats const keep = dv.workers[0].stats;
            ResidentSet probe(ctx, dv, dv.workers[0]);
            try
            {
                probe.init(nS);
                probe.admit(0, nS, dv.workers[0].stream, 0, &dv.pilot, sample.data());
            }
            catch (...) { probe.release(); throw; }
            probe.release();
            dv.workers[0].stats = keep;
        }
        CK(cudaEventRecord(dv.evRun0, dv.stream));
        // the other workers' streams start after the start event
        for (int w = 1; w < nW; ++w) CK(cudaStreamWaitEvent(dv.workers[(size_t)w].stream, dv.evRun0, 0));
        if (nW == 1)
            run_worker(ctx, dv, dv.workers[0], 1);
        else
        {
            std::vector<std::thread> th;
            for (int w = 0; w < nW; ++w)
                th.emplace_back([&ctx, &dv, w, nW]() { run_worker(ctx, dv, dv.workers[(size_t)w], nW); });
            for (std::thread & t : th) t.join();
        }
        CK(cudaSetDevice(dv.id));
        for (int w = 1; w < nW; ++w) CK(cudaStreamWaitEvent(dv.stream, dv.workers[(size_t)w].evDone, 0));
        CK(cudaEventRecord(dv.evRun1, dv.stream));
        CK(cudaStreamSynchronize(dv.stream));
        dv.stats.ms_device = ev_ms(dv.evRun0, dv.evRun1);
        for (int w = 0; w < nW; ++w)
        {
            Worker & wk = dv.workers[(size_t)w];
            if (!wk.err.empty() && dv.err.empty()) dv.err = wk.err;
            dv.stats.ms_setup += wk.stats.ms_setup; dv.stats.ms_dp += wk.stats.ms_dp; dv.stats.ms_bounds += wk.stats.ms_bounds;
            dv.stats.dp_cells += wk.stats.dp_cells; dv.stats.dp_launches += wk.stats.dp_launches;
            dv.stats.kernel_launches += wk.stats.kernel_launches; dv.stats.filter_cells += wk.stats.filter_cells;
            dv.stats.n_pairs += wk.stats.n_pairs; dv.stats.sum_iterations += wk.stats.sum_iterations;
            dv.stats.n_active_edges += wk.stats.n_active_edges; dv.stats.n_interactions += wk.stats.n_interactions;
        }
        dv.stats.reserved[0] = nW;
        for (int w = 0; w < nW; ++w)
        {
            lara_stats const & ws = dv.workers[(size_t)w].stats;
            for (int k : {1, 2, 3, 4, 5, 6, 7, 8, 9, 12}) dv.stats.reserved[k] += ws.reserved[k];
            if (ws.reserved[11] > dv.stats.reserved[11]) { dv.stats.reserved[10] = ws.reserved[10]; dv.stats.reserved[11] = ws.reserved[11]; }
        }
    }
    catch (std::exception const & e)
    {
        dv.err = e.what();
        cudaGetLastError();
    }
}

// contexts per device that rely on the raised release threshold of the device's default memory pool
std::mutex g_poolMutex;
std::map<int, int> g_poolUsers;
void pool_acquire(int dev) { std::lock_guard<std::mutex> lock(g_poolMutex); ++g_poolUsers[dev]; }
bool pool_release(int dev)
{
    std::lock_guard<std::mutex> lock(g_poolMutex);
    auto it = g_poolUsers.find(dev);
    if (it == g_poolUsers.end()) return false;
    if (--it->second > 0) return false;
    g_poolUsers.erase(it);
    return true;
}

int set_err(lara_ctx * ctx, int code, std::string const & msg)
{
    if (ctx) ctx->err = msg; else g_createErr = msg;
    return code;
}

} // namespace

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------
extern "C" {

int lara_gpu_version(void) { return LARA_ABI_VERSION; }

int lara_gpu_visible_devices(void)
{
    int count = 0;
    return cudaGetDeviceCount(&count) == cudaSuccess ? count : 0;
}

char const * lara_gpu_last_error(lara_ctx * ctx) { return ctx ? ctx->err.c_str() : g_createErr.c_str(); }

int lara_gpu_create(lara_ctx ** out, lara_params const * params, int n_devices, int const * device_ids)
{
    if (!out || !params) return set_err(nullptr, LARA_E_INVALID, "null argument");
    *out = nullptr;
    if (params->matching == 0)
        return set_err(nullptr, LARA_E_UNSUPPORTED, "-m 0 (LEMON maximum weighted matching) is outside the GPU path");
    if (params->matching > 8)
        return set_err(nullptr, LARA_E_UNSUPPORTED, "greedy matching lookahead > 8 is not supported on the GPU path");
    if (params->num_iterations < 1 || params->max_nondecreasing < 0)
        return set_err(nullptr, LARA_E_INVALID, "invalid iteration parameters");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return set_err(nullptr, LARA_E_NO_DEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                                  " (lara_b200 has no CPU fallback)");
    std::vector<int> ids;
    if (n_devices <= 0 || !device_ids)
    {
        int cur = 0;
        cudaGetDevice(&cur);
        ids.push_back(cur);
    }
    else
        ids.assign(device_ids, device_ids + n_devices);
    lara_ctx * ctx = new lara_ctx();
    if (char const * e = std::getenv("LARA_B200_WORKERS")) ctx->nWorkers = std::max(1, std::min(8, std::atoi(e)));
    if (char const * e = std::getenv("LARA_B200_BAND")) ctx->bandMode = std::atoi(e) != 0;
    if (char const * e = std::getenv("LARA_B200_MARGIN")) ctx->bandMargin = std::max(0, std::min(16, std::atoi(e)));
    if (char const * e = std::getenv("LARA_B200_BANDMAXW")) ctx->bandMaxW = std::max(0, std::min(1024, std::atoi(e)));
    if (char const * e = std::getenv("LARA_B200_BANDMINWARPS")) ctx->bandMinWarps = std::max(1, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_BAND24")) ctx->band24 = std::atoi(e) != 0;
    if (char const * e = std::getenv("LARA_B200_BANDCLASSMINWARPS")) ctx->bandClassMinWarps = std::max(1, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_BAND24MINWARPS")) ctx->band24MinWarps = std::max(1, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_BANDQ")) ctx->bandForceQ = std::atoi(e);
    if (char const * e = std::getenv("LARA_B200_VERBOSE")) ctx->verbose = std::atoi(e);
    if (char const * e = std::getenv("LARA_B200_SMEMCAND")) ctx->postSmemCand = std::max(32, std::min(2048, std::atoi(e)));
    if (char const * e = std::getenv("LARA_B200_BATCH")) ctx->batchPairsDefault = (size_t)std::max(32, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_BATCHCAP")) ctx->batchPairsCap = (size_t)std::max(0, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_POSTORDER")) ctx->postInOrder = std::atoi(e) != 0;
    if (char const * e = std::getenv("LARA_B200_SPLIT")) ctx->splitLaunch = std::atoi(e) != 0;
    if (char const * e = std::getenv("LARA_B200_SPLITWARPS")) ctx->bandSplitWarps = std::max(1, std::atoi(e));
    if (char const * e = std::getenv("LARA_B200_ROWS")) ctx->forceRows = std::atoi(e) != 0;
    if (char const * e = std::getenv("LARA_B200_REFILL")) ctx->refillMode = std::max(0, std::min(2, std::atoi(e)));
    if (char const * e = std::getenv("LARA_B200_MEMBUDGET")) ctx->memBudget = 1e6 * std::max(0.0, std::atof(e));
    if (char const * e = std::getenv("LARA_B200_REFILLSHARE")) ctx->refillFirstShare = std::max(0.1, std::min(1.0, std::atof(e)));
    if (char const * e = std::getenv("LARA_B200_WORKERPAIRS")) ctx->workerMinPairs = (size_t)std::max(1, std::atoi(e));
    ctx->P = *params;
    ctx->dp = make_dev_params(*params);
    // gap costs outside go < ge: no band kernels.  open == extend: the scalar reference gets linear gap costs from SeqAn
    // (global_alignment_unbanded.h:775-778), the SIMD build asks for affine ones (src/subgradient_solver_simd.hpp:257-263);
    // open > extend: affine
    if (params->gap_open == params->gap_extend) ctx->rowsMode = params->step_mode == LARA_STEP_SCALAR_FLOAT ? 1 : 0;
    else if (params->gap_open > params->gap_extend) ctx->rowsMode = 2;
    try
    {
        for (int id : ids)
        {
            if (id < 0 || id >= count) throw std::runtime_error("device id out of range");
            cudaDeviceProp prop;
            CK(cudaGetDeviceProperties(&prop, id));
            if (prop.major < 10)
                throw std::runtime_error(std::string("device ") + prop.name + " is not sm_100 (built for sm_100a only)");
            ctx->devs.emplace_back();
            Device & dv = ctx->devs.back();
            dv.id = id;
            CK(cudaSetDevice(id));
            for (int w = 0; w < ctx->nWorkers; ++w)
            {
                dv.workers.emplace_back();
                Worker & wk = dv.workers.back();
                CK(cudaStreamCreateWithFlags(&wk.stream, cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&wk.evDone, cudaEventDisableTiming));
                for (int k = 0; k < 2; ++k)
                {
                    cudaEvent_t ev;
                    CK(cudaEventCreate(&ev));
                    wk.events.push_back(ev);
                }
            }
            dv.stream = dv.workers[0].stream;
            CK(cudaMalloc((void **)&dv.negBuf, 64 * sizeof(int32_t)));
            std::vector<int32_t> neg(64, kDpNeg);
            CK(cudaMemcpy(dv.negBuf, neg.data(), 64 * sizeof(int32_t), cudaMemcpyHostToDevice));
            for (int k = 0; k < 2; ++k)
            {
                cudaEvent_t ev;
                CK(cudaEventCreate(&ev));
                dv.events.push_back(ev);
            }
            CK(cudaEventCreate(&dv.evRun0));
            CK(cudaEventCreate(&dv.evRun1));
            CK(post_configure_device());
            CK(cudaDeviceGetAttribute(&dv.smCount, cudaDevAttrMultiProcessorCount, id));
            CK(cudaDeviceGetAttribute(&dv.smemOptin, cudaDevAttrMaxSharedMemoryPerBlockOptin, id));
            // L2 fetch granularity (a hint; the driver default fetches 64 bytes per missed 32-byte sector): the post
            // kernel reads isolated 32-byte records, the fill reads whole lines either way.  LARA_B200_L2FETCH = 32|64|128
            if (char const * e = std::getenv("LARA_B200_L2FETCH"))
            {
                int const g = std::atoi(e);
                if (g == 32 || g == 64 || g == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)g);
                size_t got = 0;
                cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
                if (ctx->verbose) std::fprintf(stderr, "[lara_b200] dev %d L2 fetch granularity %zu\n", id, got);
            }
            // keep freed blocks in the stream-ordered pool between batches (undone by the last lara_gpu_destroy)
            cudaMemPool_t pool;
            CK(cudaDeviceGetDefaultMemPool(&pool, id));
            uint64_t thr = UINT64_MAX;
            CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
            pool_acquire(id);
            dv.poolHeld = true;
        }
    }
    catch (std::exception const & ex)
    {
        g_createErr = ex.what();
        lara_gpu_destroy(ctx);
        return LARA_E_CUDA;
    }
    *out = ctx;
    return LARA_OK;
}

void lara_gpu_destroy(lara_ctx * ctx)
{
    if (!ctx) return;
    for (Device & dv : ctx->devs)
    {
        cudaSetDevice(dv.id);
        for (Worker & wk : dv.workers) if (wk.stream) cudaStreamSynchronize(wk.stream);
        if (dv.poolHeld && pool_release(dv.id))
        {
            // the stream-ordered pool kept freed blocks between batches (lara_gpu_create); this was the last context
            // on the device: give them back and restore the default threshold
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dv.id) == cudaSuccess)
            {
                uint64_t thr = 0;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
                cudaMemPoolTrimTo(pool, 0);
            }
            cudaGetLastError();
        }
        device_free_inputs(dv);
        device_free_outputs(dv);
        cudaFree(dv.negBuf);
        for (cudaEvent_t ev : dv.events) cudaEventDestroy(ev);
        if (dv.evRun0) cudaEventDestroy(dv.evRun0);
        if (dv.evRun1) cudaEventDestroy(dv.evRun1);
        for (Worker & wk : dv.workers)
        {
            if (wk.hFinished) cudaFreeHost(wk.hFinished);
            if (wk.hTable) cudaFreeHost(wk.hTable);
            if (wk.evTable) cudaEventDestroy(wk.evTable);
            for (int k = 0; k < 2; ++k) if (wk.evBlock[k]) cudaEventDestroy(wk.evBlock[k]);
            for (cudaEvent_t ev : wk.events) cudaEventDestroy(ev);
            if (wk.evDone) cudaEventDestroy(wk.evDone);
            if (wk.evFork) cudaEventDestroy(wk.evFork);
            if (wk.evSide) cudaEventDestroy(wk.evSide);
            if (wk.side) cudaStreamDestroy(wk.side);
            cudaFree(wk.slab);
            for (cudaEvent_t ev : wk.auxJoin) cudaEventDestroy(ev);
            for (cudaStream_t st : wk.aux) cudaStreamDestroy(st);
            if (wk.stream) cudaStreamDestroy(wk.stream);
        }
    }
    delete ctx;
}

int lara_gpu_set_sequences(lara_ctx * ctx, int32_t n_seq, uint8_t const * seq_cat, int64_t const * seq_off,
                           int32_t const * bp_partner, int32_t const * bp_weight, int64_t const * bp_off)
{
    if (!ctx) return LARA_E_INVALID;
    if (n_seq < 0 || !seq_off || !bp_off) return set_err(ctx, LARA_E_INVALID, "null argument");
    int64_t const total = seq_off[n_seq];
    int64_t const nnz = bp_off[total];
    if (total > 0 && !seq_cat) return set_err(ctx, LARA_E_INVALID, "null argument: seq_cat");
    if (nnz > 0 && (!bp_partner || !bp_weight)) return set_err(ctx, LARA_E_INVALID, "null argument: bp_partner / bp_weight");
    for (int32_t i = 0; i < n_seq; ++i)
    {
        int64_t const L = seq_off[i + 1] - seq_off[i];
        if (L < 1) return set_err(ctx, LARA_E_INVALID, "empty sequence");
        if (L > 32767) return set_err(ctx, LARA_E_UNSUPPORTED, "sequences longer than 32767 nt are not supported");
        for (int64_t g = seq_off[i]; g < seq_off[i + 1]; ++g)
        {
            if (seq_cat[g] > 4) return set_err(ctx, LARA_E_INVALID, "sequence code outside 0..4");
            for (int64_t x = bp_off[g]; x < bp_off[g + 1]; ++x)
                if (bp_partner[x] < 0 || bp_partner[x] >= L) return set_err(ctx, LARA_E_INVALID, "base pair partner out of range");
        }
        // the adjacency is that of an undirected graph (src/lagrange.hpp:130-138): every base pair listed once from
        // each end with the same weight; the candidate buffer of the matching is sized from that
        for (int64_t g = seq_off[i]; g < seq_off[i + 1]; ++g)
            for (int64_t x = bp_off[g]; x < bp_off[g + 1]; ++x)
            {
                int64_t const h = seq_off[i] + bp_partner[x];
                int back = 0;
                for (int64_t y = bp_off[h]; y < bp_off[h + 1]; ++y)
                    if (seq_off[i] + bp_partner[y] == g) { ++back; if (bp_weight[y] != bp_weight[x]) back = -1000; }
                if (h == g || back != 1)
                    return set_err(ctx, LARA_E_INVALID, "base pair adjacency is not symmetric and duplicate-free (sequence " +
                                                        std::to_string(i) + ", position " + std::to_string(g - seq_off[i]) + ")");
            }
    }
    ctx->nSeq = n_seq;
    ctx->seqOff.assign(seq_off, seq_off + n_seq + 1);
    ctx->bpOff.assign(bp_off, bp_off + total + 1);
    ctx->nbp.assign((size_t)n_seq, 0);
    for (int32_t i = 0; i < n_seq; ++i)
        ctx->nbp[(size_t)i] = (int32_t)((bp_off[seq_off[i + 1]] - bp_off[seq_off[i]] + 1) / 2);
    ctx->prepared = false;
    try
    {
        for (Device & dv : ctx->devs)
        {
            CK(cudaSetDevice(dv.id));
            cudaEvent_t e0 = dv.events[0], e1 = dv.events[1];
            CK(cudaEventRecord(e0, dv.stream));
            ensure_device(dv.dSeq, dv.capSeq, (size_t)std::max<int64_t>(total, 1));
            ensure_device(dv.dSeqOff, dv.capSeqOff, sizeof(int64_t) * (size_t)(n_seq + 1));
            ensure_device(dv.dBpOff, dv.capBpOff, sizeof(int64_t) * (size_t)(total + 1));
            ensure_device(dv.dBpPartner, dv.capBpPartner, sizeof(int32_t) * (size_t)std::max<int64_t>(nnz, 1));
            ensure_device(dv.dBpWeight, dv.capBpWeight, sizeof(int32_t) * (size_t)std::max<int64_t>(nnz, 1));
            CK(cudaMemcpyAsync(dv.dSeq, seq_cat, (size_t)total, cudaMemcpyHostToDevice, dv.stream));
            CK(cudaMemcpyAsync(dv.dSeqOff, seq_off, sizeof(int64_t) * (size_t)(n_seq + 1), cudaMemcpyHostToDevice, dv.stream));
            CK(cudaMemcpyAsync(dv.dBpOff, bp_off, sizeof(int64_t) * (size_t)(total + 1), cudaMemcpyHostToDevice, dv.stream));
            if (nnz)
            {
                CK(cudaMemcpyAsync(dv.dBpPartner, bp_partner, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, dv.stream));
                CK(cudaMemcpyAsync(dv.dBpWeight, bp_weight, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, dv.stream));
            }
            CK(cudaEventRecord(e1, dv.stream));
            CK(cudaStreamSynchronize(dv.stream));
            dv.in.seqCat = (uint8_t const *)dv.dSeq; dv.in.seqOff = (int64_t const *)dv.dSeqOff;
            dv.in.bpOff = (int64_t const *)dv.dBpOff; dv.in.bpPartner = (int32_t const *)dv.dBpPartner;
            dv.in.bpWeight = (int32_t const *)dv.dBpWeight;
            ctx->stats.ms_h2d = ev_ms(e0, e1);
        }
        ctx->stats.h2d_bytes = total + 8 * (n_seq + 1) + 8 * (total + 1) + 8 * nnz;
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

int lara_gpu_prepare_pairs(lara_ctx * ctx, int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b)
{
    if (!ctx) return LARA_E_INVALID;
    if (n_pairs < 0 || (n_pairs > 0 && (!idx_a || !idx_b))) return set_err(ctx, LARA_E_INVALID, "null argument");
    if (ctx->seqOff.empty()) return set_err(ctx, LARA_E_INVALID, "lara_gpu_set_sequences must be called first");
    std::vector<OrientedPair> all((size_t)n_pairs);
    for (int64_t p = 0; p < n_pairs; ++p)
    {
        int32_t const i = std::min(idx_a[p], idx_b[p]), j = std::max(idx_a[p], idx_b[p]);
        if (i < 0 || j >= ctx->nSeq || i == j) return set_err(ctx, LARA_E_INVALID, "pair index out of range");
        int32_t const li = (int32_t)(ctx->seqOff[i + 1] - ctx->seqOff[i]), lj = (int32_t)(ctx->seqOff[j + 1] - ctx->seqOff[j]);
        OrientedPair & op = all[(size_t)p];
        if (li >= lj) { op.first = i; op.second = j; op.la = li; op.lb = lj; }       // solver :117-120
        else { op.first = j; op.second = i; op.la = lj; op.lb = li; }
        op.request = p;
        if (op.la > 32767)
            return set_err(ctx, LARA_E_UNSUPPORTED, "sequences longer than 32,767 nt are not supported (16-bit positions)");
    }
    {
        // the post kernel keeps the lines of a pair in shared memory (kernels_post.cu: WarpSmem)
        int maxLa = 0, maxLb = 0;
        for (OrientedPair const & op : all) { maxLa = std::max(maxLa, op.la); maxLb = std::max(maxLb, op.lb); }
        for (Device const & dv : ctx->devs)
            if (n_pairs > 0 && post_smem_bytes_per_warp(maxLa, maxLb, 32) > (size_t)dv.smemOptin)
                return set_err(ctx, LARA_E_UNSUPPORTED, "a pair of " + std::to_string(maxLa) + " x " + std::to_string(maxLb) +
                               " nt needs " + std::to_string(post_smem_bytes_per_warp(maxLa, maxLb, 32)) +
                               " bytes of shared memory per warp in the iteration kernel, the device offers " +
                               std::to_string(dv.smemOptin));
    }
    // longest first, like the reference's queue (CompareSeqLength, src/io.hpp:352-365)
    std::stable_sort(all.begin(), all.end(), [](OrientedPair const & x, OrientedPair const & y) {
        return x.la != y.la ? x.la > y.la : x.lb > y.lb;
    });
    try
    {
        size_t const nd = ctx->devs.size();
        for (size_t k = 0; k < nd; ++k)
        {
            Device & dv = ctx->devs[k];
            CK(cudaSetDevice(dv.id));
            dv.pairs.clear(); dv.pilot = Pilot{};
            for (size_t q = k; q < all.size(); q += nd) dv.pairs.push_back(all[q]);   // interleaved shares
            dv.outOff.assign(dv.pairs.size() + 1, 0);
            for (size_t q = 0; q < dv.pairs.size(); ++q) dv.outOff[q + 1] = dv.outOff[q] + 3ll * dv.pairs[q].lb;
            dv.outLinesTotal = dv.outOff.back();
            size_t const np = std::max<size_t>(dv.pairs.size(), 1);
            ensure_device(dv.dOutHdr, dv.capOutHdr, sizeof(PairOutHdr) * np);
            ensure_device(dv.dOutLines, dv.capOutLines, sizeof(int32_t) * (size_t)std::max<int64_t>(dv.outLinesTotal, 1));
            ensure_device(dv.dOutOff, dv.capOutOff, sizeof(int64_t) * (np + 1));
            CK(cudaMemcpy(dv.dOutOff, dv.outOff.data(), sizeof(int64_t) * dv.outOff.size(), cudaMemcpyHostToDevice));
        }
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    ctx->nPrepared = n_pairs;
    ctx->prepared = true;
    ctx->ran = false;
    return LARA_OK;
}

int lara_gpu_run_prepared(lara_ctx * ctx)
{
    if (!ctx) return LARA_E_INVALID;
    if (!ctx->prepared) return set_err(ctx, LARA_E_INVALID, "no prepared pairs");
    ctx->ran = false;
    auto const t0 = std::chrono::steady_clock::now();
    if (ctx->devs.size() == 1)
        run_device(*ctx, ctx->devs[0]);
    else
    {
        std::vector<std::thread> th;
        for (Device & dv : ctx->devs) th.emplace_back([ctx, &dv]() { run_device(*ctx, dv); });
        for (std::thread & t : th) t.join();
    }
    double const ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    lara_stats agg{};
    agg.ms_h2d = ctx->stats.ms_h2d; agg.h2d_bytes = ctx->stats.h2d_bytes;
    {
        std::string all;
        for (Device & dv : ctx->devs)
            if (!dv.err.empty()) { all += (all.empty() ? "" : "; ") + ("device " + std::to_string(dv.id) + ": " + dv.err); dv.err.clear(); }
        if (!all.empty()) return set_err(ctx, LARA_E_CUDA, all);
    }
    for (Device & dv : ctx->devs)
    {
        agg.ms_setup = std::max(agg.ms_setup, dv.stats.ms_setup);
        agg.ms_dp = std::max(agg.ms_dp, dv.stats.ms_dp);
        agg.ms_bounds = std::max(agg.ms_bounds, dv.stats.ms_bounds);
        agg.ms_device = std::max(agg.ms_device, dv.stats.ms_device);
        agg.dp_cells += dv.stats.dp_cells; agg.dp_launches += dv.stats.dp_launches;
        agg.kernel_launches += dv.stats.kernel_launches; agg.filter_cells += dv.stats.filter_cells;
        agg.n_pairs += dv.stats.n_pairs; agg.sum_iterations += dv.stats.sum_iterations;
        agg.n_active_edges += dv.stats.n_active_edges; agg.n_interactions += dv.stats.n_interactions;
        agg.reserved[0] = std::max(agg.reserved[0], dv.stats.reserved[0]);     // concurrent workers per device
        // [1] pairs solved by the band kernels, [2] ... in blocks of 16 diagonals, [3] sum of the band widths, [4] algorithmic
        // HBM bytes of the DP launches, [5] DP cells actually computed, [6] cohorts, [7] iteration steps, [8] slots of
        // the steps, [9] live slots of the steps, [12] pairs taken into running sets
        for (int k : {1, 2, 3, 4, 5, 6, 7, 8, 9, 12}) agg.reserved[k] += dv.stats.reserved[k];
        if (dv.stats.reserved[11] > agg.reserved[11]) { agg.reserved[10] = dv.stats.reserved[10]; agg.reserved[11] = dv.stats.reserved[11]; }
    }
    agg.ms_total = ms;
    ctx->stats = agg;
    ctx->ran = true;
    return LARA_OK;
}

int lara_gpu_fetch_prepared(lara_ctx * ctx, lara_results * out)
{
    if (!ctx || !out) return LARA_E_INVALID;
    if (!ctx->ran) return set_err(ctx, LARA_E_INVALID, "lara_gpu_run_prepared has not run");
    ResultOwner * own = new ResultOwner();
    own->pairs.resize((size_t)ctx->nPrepared);
    try
    {
        // 1. headers of all devices -> number of lines per pair, position of each pair's lines in the result arrays
        //    (device by device, in the device's own pair order; lara_pair_result::lines_off says where)
        std::vector<std::vector<PairOutHdr>> hdr(ctx->devs.size());
        std::vector<std::vector<int64_t>> dstOff(ctx->devs.size());
        std::vector<int64_t> devBase(ctx->devs.size() + 1, 0);
        int64_t d2h = 0;
        for (size_t k = 0; k < ctx->devs.size(); ++k)
        {
            Device & dv = ctx->devs[k];
            CK(cudaSetDevice(dv.id));
            hdr[k].resize(dv.pairs.size());
            if (!dv.pairs.empty())
                CK(cudaMemcpyAsync(hdr[k].data(), dv.dOutHdr, sizeof(PairOutHdr) * dv.pairs.size(), cudaMemcpyDeviceToHost, dv.stream));
            CK(cudaStreamSynchronize(dv.stream));
            dstOff[k].resize(dv.pairs.size() + 1);
            int64_t run = 0;
            for (size_t q = 0; q < dv.pairs.size(); ++q) { dstOff[k][q] = run; run += hdr[k][q].nLines; }
            dstOff[k][dv.pairs.size()] = run;
            devBase[k + 1] = devBase[k] + run;
            d2h += (int64_t)(sizeof(PairOutHdr) * dv.pairs.size());
        }
        int64_t const total = devBase.back();
        own->nLines = total;
        own->allocate(total);
        // 2. gather on the device, three dense copies per device
        double msD2h = 0;
        std::vector<int32_t *> scratch(ctx->devs.size(), nullptr);
        for (size_t k = 0; k < ctx->devs.size(); ++k)
        {
            Device & dv = ctx->devs[k];
            int64_t const nL = dstOff[k].back();
            if (dv.pairs.empty() || nL == 0) continue;
            CK(cudaSetDevice(dv.id));
            int64_t * dDst = nullptr;
            CK(cudaMallocAsync((void **)&scratch[k], sizeof(int32_t) * 3 * (size_t)nL, dv.stream));
            CK(cudaMallocAsync((void **)&dDst, sizeof(int64_t) * dstOff[k].size(), dv.stream));
            CK(cudaMemcpyAsync(dDst, dstOff[k].data(), sizeof(int64_t) * dstOff[k].size(), cudaMemcpyHostToDevice, dv.stream));
            CK(cudaEventRecord(dv.events[0], dv.stream));
            CK(launch_pack_lines((int)dv.pairs.size(), dv.dOutHdr, dv.dOutLines, dv.dOutOff, dDst, scratch[k], scratch[k] + nL,
                                 scratch[k] + 2 * nL, dv.stream));
            CK(cudaMemcpyAsync(own->posLow + devBase[k], scratch[k], sizeof(int32_t) * (size_t)nL, cudaMemcpyDeviceToHost, dv.stream));
            CK(cudaMemcpyAsync(own->posHigh + devBase[k], scratch[k] + nL, sizeof(int32_t) * (size_t)nL, cudaMemcpyDeviceToHost, dv.stream));
            CK(cudaMemcpyAsync(own->weight + devBase[k], scratch[k] + 2 * nL, sizeof(int32_t) * (size_t)nL, cudaMemcpyDeviceToHost, dv.stream));
            CK(cudaEventRecord(dv.events[1], dv.stream));
            CK(cudaFreeAsync(dDst, dv.stream));
            CK(cudaFreeAsync(scratch[k], dv.stream));
            d2h += 12 * nL;
        }
        // 3. pair records in request order (while the copies run)
        for (size_t k = 0; k < ctx->devs.size(); ++k)
        {
            Device & dv = ctx->devs[k];
            for (size_t q = 0; q < dv.pairs.size(); ++q)
            {
                OrientedPair const & op = dv.pairs[q];
                PairOutHdr const & h = hdr[k][q];
                lara_pair_result & r = own->pairs[(size_t)op.request];
                r.idx_low = std::min(op.first, op.second);
                r.idx_high = std::max(op.first, op.second);
                r.best_lower = h.bestLower; r.best_upper = h.bestUpper; r.iterations = h.iterations; r.n_lines = h.nLines;
                r.lines_off = devBase[k] + dstOff[k][q];
            }
        }
        for (size_t k = 0; k < ctx->devs.size(); ++k)
        {
            Device & dv = ctx->devs[k];
            if (dv.pairs.empty() || dstOff[k].back() == 0) continue;
            CK(cudaSetDevice(dv.id));
            CK(cudaStreamSynchronize(dv.stream));
            msD2h = std::max(msD2h, ev_ms(dv.events[0], dv.events[1]));
        }
        ctx->stats.ms_d2h = msD2h;
        ctx->stats.d2h_bytes = d2h;
    }
    catch (std::exception const & ex)
    {
        // device-to-host copies into the block may still be in flight
        for (Device & dv : ctx->devs) { cudaSetDevice(dv.id); cudaStreamSynchronize(dv.stream); }
        cudaGetLastError();
        delete own;
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    out->n_pairs = ctx->nPrepared;
    out->pairs = own->pairs.data();
    out->n_lines_total = own->nLines;
    out->line_pos_low = own->posLow;
    out->line_pos_high = own->posHigh;
    out->line_weight = own->weight;
    out->owner_ = own;
    return LARA_OK;
}

int lara_gpu_solve_pairs(lara_ctx * ctx, int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b, lara_results * out)
{
    int rc = lara_gpu_prepare_pairs(ctx, n_pairs, idx_a, idx_b);
    if (rc != LARA_OK) return rc;
    rc = lara_gpu_run_prepared(ctx);
    if (rc != LARA_OK) return rc;
    return lara_gpu_fetch_prepared(ctx, out);
}

int lara_gpu_solve_all_pairs(lara_ctx * ctx, lara_results * out)
{
    if (!ctx) return LARA_E_INVALID;
    std::vector<int32_t> a, b;
    for (int32_t i = 0; i + 1 < ctx->nSeq; ++i)
        for (int32_t j = i + 1; j < ctx->nSeq; ++j) { a.push_back(i); b.push_back(j); }
    return lara_gpu_solve_pairs(ctx, (int64_t)a.size(), a.data(), b.data(), out);
}

void lara_gpu_free_results(lara_results * out)
{
    if (out && out->owner_)
    {
        delete static_cast<ResultOwner *>(out->owner_);
        std::memset(out, 0, sizeof(*out));
    }
}

int lara_gpu_get_stats(lara_ctx * ctx, lara_stats * stats)
{
    if (!ctx || !stats) return LARA_E_INVALID;
    *stats = ctx->stats;
    return LARA_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------
// unit-level entry points (parity tests of single kernels)
// ---------------------------------------------------------------------------------------------------
extern "C" {

int lara_gpu_gotoh_batch(lara_ctx * ctx, int32_t n_problems, int32_t lenA, int32_t lenB, int32_t const * pssm,
                         int32_t gap_open, int32_t gap_extend, int32_t * score, int32_t * n_lines, int32_t * lines_a,
                         int32_t * lines_b, int32_t * gap_score)
{
    if (!ctx || !pssm || n_problems < 1 || lenA < 1 || lenB < 1) return set_err(ctx, LARA_E_INVALID, "bad argument");
    if ((lenB + 31) / 32 > kMaxR) return set_err(ctx, LARA_E_UNSUPPORTED, "lenB exceeds the single-pass DP limit");
    Device & dv = ctx->devs[0];
    cudaStream_t const s = dv.stream;
    int const n = n_problems;
    try
    {
        CK(cudaSetDevice(dv.id));
        std::vector<PairDesc> desc((size_t)n);
        Region rg;
        for (int pass = 0; pass < 2; ++pass)
        {
            rg.used = 0;
            for (int i = 0; i < n; ++i)
            {
                PairDesc & d = desc[(size_t)i];
                std::memset(&d, 0, sizeof(d));
                d.la = lenA; d.lb = lenB; d.R = (lenB + 31) / 32; d.W = d.R <= 8 ? 1 : (d.R <= 16 ? 2 : 4);
                d.nSteps = (lenA + 31 + 3) & ~3; d.bitWords = (lenB + 31) / 32;
                d.nVals = 32 * lenA * d.R;
                d.offFlags = rg.take(32ll * ((d.nSteps + 31) / 32));
                d.offLanePtr = rg.take(32);
                d.offVals = rg.take(d.nVals);
                d.offTrace = rg.take(32ll * d.nSteps * d.W);
                d.offLines = rg.take(lenB);
            }
            if (pass == 0) CK(cudaMalloc((void **)&rg.ptr, (size_t)rg.used * 4));
        }
        int32_t * arena = rg.ptr;
        PairDesc * dDesc; PairState * dState; int32_t * dPssm;
        size_t const nP = (size_t)n * lenA * lenB;
        CK(cudaMalloc((void **)&dDesc, sizeof(PairDesc) * (size_t)n));
        CK(cudaMalloc((void **)&dState, sizeof(PairState) * (size_t)n));
        CK(cudaMalloc((void **)&dPssm, sizeof(int32_t) * nP));
        CK(cudaMemcpyAsync(dDesc, desc.data(), sizeof(PairDesc) * (size_t)n, cudaMemcpyHostToDevice, s));
        CK(cudaMemsetAsync(dState, 0, sizeof(PairState) * (size_t)n, s));
        CK(cudaMemcpyAsync(dPssm, pssm, sizeof(int32_t) * nP, cudaMemcpyHostToDevice, s));
        CK(launch_dense_to_stream(n, dDesc, arena, dPssm, s));
        CK(launch_gotoh_fill(desc[0].R, n, nullptr, dDesc, dState, arena, dv.negBuf, gap_open, gap_extend, s));
        CK(launch_traceback_only(n, dDesc, dState, arena, gap_open, gap_extend, lenA, lenB, s));
        std::vector<PairState> hs((size_t)n);
        CK(cudaMemcpyAsync(hs.data(), dState, sizeof(PairState) * (size_t)n, cudaMemcpyDeviceToHost, s));
        std::vector<int32_t> lines((size_t)n * lenB);
        for (int i = 0; i < n; ++i)
            CK(cudaMemcpyAsync(lines.data() + (size_t)i * lenB, arena + desc[(size_t)i].offLines, sizeof(int32_t) * (size_t)lenB,
                               cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        for (int i = 0; i < n; ++i)
        {
            if (score) score[i] = hs[(size_t)i].currentUpper;
            if (n_lines) n_lines[i] = hs[(size_t)i].nLines;
            if (gap_score) gap_score[i] = hs[(size_t)i].gapScore;
            for (int k = 0; k < hs[(size_t)i].nLines; ++k)
            {
                uint32_t const v = (uint32_t)lines[(size_t)i * lenB + k];
                if (lines_a) lines_a[(size_t)i * lenB + k] = (int32_t)(v >> 16);
                if (lines_b) lines_b[(size_t)i * lenB + k] = (int32_t)(v & 0xffffu);
            }
        }
        cudaFree(dPssm); cudaFree(dState); cudaFree(dDesc); cudaFree(arena);
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

int lara_gpu_gotoh_band_batch(lara_ctx * ctx, int32_t n_problems, int32_t lenA, int32_t lenB, int32_t const * pssm,
                              int32_t d_lo, int32_t band_width, int32_t lanes, int32_t gap_open, int32_t gap_extend,
                              int32_t * score, int32_t * n_lines, int32_t * lines_a, int32_t * lines_b, int32_t * gap_score)
{
    if (!ctx || !pssm || n_problems < 1 || lenA < 1 || lenB < 1) return set_err(ctx, LARA_E_INVALID, "bad argument");
    int const G = lanes, W = band_width, Q = (G > 0 && W % (2 * G) == 0) ? W / (2 * G) : 0;     // (24, 1) -> Q = 12
    if (!band_shape_supported(G, Q) || d_lo > 0 || d_lo + W - 1 < 0)
        return set_err(ctx, LARA_E_INVALID, "band: width / (2 * lanes) must be 16 (lanes 1..32), 8 (lanes 2..32) or 4 (lanes 4..32), "
                                            "or 12 (width 24, one lane), and d_lo <= 0 <= d_lo + width - 1");
    if (!(gap_open < gap_extend)) return set_err(ctx, LARA_E_UNSUPPORTED, "band kernel needs gap_open < gap_extend");
    Device & dv = ctx->devs[0];
    cudaStream_t const s = dv.stream;
    int const n = n_problems, dHi = d_lo + W - 1, ppw = 32 / G;
    int const nGroups = (n + ppw - 1) / ppw;
    std::vector<int32_t> slots((size_t)nGroups * ppw, -1);
    try
    {
        CK(cudaSetDevice(dv.id));
        std::vector<PairDesc> desc((size_t)n);
        Region rg;
        for (int pass = 0; pass < 2; ++pass)
        {
            rg.used = 0;
            for (int gi = 0; gi < nGroups; ++gi)
            {
                int64_t const ov = rg.take((int64_t)(lenB + 1) * W * ppw);
                if (ppw == 32) rg.used = (rg.used + 7) & ~(int64_t)7;
                int64_t const ot = ppw == 32 ? rg.take((int64_t)((lenB + 2) & ~1) * 4 * 32)
                                             : rg.take((int64_t)(lenB + 1) * (W / 8) * ppw);
                int64_t const ou = rg.take(4ll * ppw * (lenB + 1));
                int64_t const ol = rg.take(4ll * ppw * (lenA + 1));
                for (int j = 0; j < ppw && gi * ppw + j < n; ++j)
                {
                    int const i = gi * ppw + j;
                    PairDesc & d = desc[(size_t)i];
                    std::memset(&d, 0, sizeof(d));
                    d.la = lenA; d.lb = lenB; d.bandW = W; d.bandG = G; d.bandQ = Q; d.bandSlot = j; d.dLo = d_lo; d.dHi = dHi;
                    d.offBVals = ov; d.offBTrace = ot;
                    d.offU = ou; d.offL = ol;
                    d.offLines = rg.take(lenB);
                    slots[(size_t)i] = i;
                }
            }
            if (pass == 0) CK(cudaMalloc((void **)&rg.ptr, (size_t)rg.used * 4));
        }
        // dense band array on the host: cell (a, b) -> row b + 1, diagonal index a - b - d_lo
        std::vector<int32_t> host((size_t)rg.used, kDpNeg);
        for (int i = 0; i < n; ++i)
            for (int a = 0; a < lenA; ++a)
                for (int b = 0; b < lenB; ++b)
                {
                    int32_t const v = pssm[((size_t)i * lenA + a) * lenB + b];
                    if (v <= kDpNeg) continue;
                    if (a - b < d_lo || a - b > dHi)
                    {
                        cudaFree(rg.ptr);
                        return set_err(ctx, LARA_E_INVALID, "band: active cell outside the band");
                    }
                    PairDesc const & d = desc[(size_t)i];
                    host[(size_t)(d.offBVals + band_val_index(W, ppw, d.bandSlot, b + 1, a - b - d_lo))] = v;
                }
        int32_t * arena = rg.ptr;
        PairDesc * dDesc; PairState * dState;
        CK(cudaMalloc((void **)&dDesc, sizeof(PairDesc) * (size_t)n));
        CK(cudaMalloc((void **)&dState, sizeof(PairState) * (size_t)n));
        CK(cudaMemcpyAsync(arena, host.data(), host.size() * 4, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(dDesc, desc.data(), sizeof(PairDesc) * (size_t)n, cudaMemcpyHostToDevice, s));
        CK(cudaMemsetAsync(dState, 0, sizeof(PairState) * (size_t)n, s));
        int32_t * dSlots;
        CK(cudaMalloc((void **)&dSlots, slots.size() * 4));
        CK(cudaMemcpyAsync(dSlots, slots.data(), slots.size() * 4, cudaMemcpyHostToDevice, s));
        CK(launch_gotoh_band(G, Q, (int)slots.size(), dSlots, dDesc, dState, arena, gap_open, gap_extend, s));
        CK(launch_traceback_only(n, dDesc, dState, arena, gap_open, gap_extend, lenA, lenB, s));
        std::vector<PairState> hs((size_t)n);
        CK(cudaMemcpyAsync(hs.data(), dState, sizeof(PairState) * (size_t)n, cudaMemcpyDeviceToHost, s));
        std::vector<int32_t> lines((size_t)n * lenB);
        for (int i = 0; i < n; ++i)
            CK(cudaMemcpyAsync(lines.data() + (size_t)i * lenB, arena + desc[(size_t)i].offLines, sizeof(int32_t) * (size_t)lenB,
                               cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        for (int i = 0; i < n; ++i)
        {
            if (score) score[i] = hs[(size_t)i].currentUpper;
            if (n_lines) n_lines[i] = hs[(size_t)i].nLines;
            if (gap_score) gap_score[i] = hs[(size_t)i].gapScore;
            for (int k = 0; k < hs[(size_t)i].nLines; ++k)
            {
                uint32_t const v = (uint32_t)lines[(size_t)i * lenB + k];
                if (lines_a) lines_a[(size_t)i * lenB + k] = (int32_t)(v >> 16);
                if (lines_b) lines_b[(size_t)i * lenB + k] = (int32_t)(v & 0xffffu);
            }
        }
        cudaFree(dSlots); cudaFree(dState); cudaFree(dDesc); cudaFree(arena);
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

int lara_gpu_edge_filter(lara_ctx * ctx, uint8_t const * seqA, int32_t lenA, uint8_t const * seqB, int32_t lenB,
                         uint8_t * active, int32_t * opt_score)
{
    if (!ctx || !seqA || !seqB || lenA < 1 || lenB < 1 || !active) return set_err(ctx, LARA_E_INVALID, "bad argument");
    Device & dv = ctx->devs[0];
    cudaStream_t const s = dv.stream;
    try
    {
        CK(cudaSetDevice(dv.id));
        // a private two-sequence input set
        std::vector<uint8_t> cat(seqA, seqA + lenA);
        cat.insert(cat.end(), seqB, seqB + lenB);
        int64_t off[3] = {0, lenA, (int64_t)lenA + lenB};
        uint8_t * dCat; int64_t * dOff;
        CK(cudaMalloc((void **)&dCat, cat.size()));
        CK(cudaMalloc((void **)&dOff, sizeof(off)));
        CK(cudaMemcpyAsync(dCat, cat.data(), cat.size(), cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(dOff, off, sizeof(off), cudaMemcpyHostToDevice, s));
        DevInputs in{};
        in.seqCat = dCat; in.seqOff = dOff;
        PairDesc d{};
        d.la = lenA; d.lb = lenB; d.seqFirst = 0; d.seqSecond = 1; d.bitWords = (lenB + 31) / 32;
        Region rg;
        d.offBits = rg.take((int64_t)lenA * d.bitWords);
        d.offScratch = rg.take((int64_t)lenA * lenB + 6ll * (lenA + 1));
        CK(cudaMalloc((void **)&rg.ptr, (size_t)rg.used * 4));
        PairDesc * dDesc; PairState * dState;
        CK(cudaMalloc((void **)&dDesc, sizeof(PairDesc)));
        CK(cudaMalloc((void **)&dState, sizeof(PairState)));
        CK(cudaMemcpyAsync(dDesc, &d, sizeof(d), cudaMemcpyHostToDevice, s));
        CK(cudaMemsetAsync(dState, 0, sizeof(PairState), s));
        CK(launch_edge_filter(1, dDesc, dState, rg.ptr, in, ctx->dp, lenB, s));
        std::vector<uint32_t> bits((size_t)lenA * d.bitWords);
        PairState hs;
        CK(cudaMemcpyAsync(bits.data(), rg.ptr + d.offBits, bits.size() * 4, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(&hs, dState, sizeof(hs), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        for (int a = 0; a < lenA; ++a)
            for (int b = 0; b < lenB; ++b)
                active[(size_t)a * lenB + b] = (uint8_t)((bits[(size_t)a * d.bitWords + (b >> 5)] >> (b & 31)) & 1u);
        if (opt_score) *opt_score = hs.optScore;
        cudaFree(dState); cudaFree(dDesc); cudaFree(rg.ptr); cudaFree(dOff); cudaFree(dCat);
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

int lara_gpu_int_issue_rate(lara_ctx * ctx, int kind, double * gops)
{
    if (!ctx || !gops) return LARA_E_INVALID;
    Device & dv = ctx->devs[0];
    try
    {
        CK(cudaSetDevice(dv.id));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, dv.id));
        int32_t * dOut;
        CK(cudaMalloc((void **)&dOut, 4));
        int const blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
        CK(launch_int_issue(kind, blocks, threads, 64, dOut, dv.stream));      // warm-up
        CK(cudaEventRecord(dv.events[0], dv.stream));
        CK(launch_int_issue(kind, blocks, threads, iters, dOut, dv.stream));
        CK(cudaEventRecord(dv.events[1], dv.stream));
        CK(cudaStreamSynchronize(dv.stream));
        double const ms = ev_ms(dv.events[0], dv.events[1]);
        double const perOp = kind == 0 ? 1.0 : (kind == 1 ? 2.0 : 5.0);
        double const ops = (double)blocks * threads * (double)iters * 32.0 * perOp;
        *gops = ops / (ms * 1e-3) / 1e9;
        cudaFree(dOut);
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------
// Entry point that allocates nothing (no counterpart in the reference; see include/lara_b200.h)
// ---------------------------------------------------------------------------------------------------
namespace
{

// Everything lara_gpu_edge_filter() takes from cudaMalloc, in one array that every device which loads the library
// holds until the process exits: pair record, solver state, sequence offsets, the two sequences, the active-edge
// bitmap and the edge filter's scratch (forward prefix scores + row buffers).  Shared by all calls: a call overwrites
// what the call before it left, and two calls that overlap on a device race.
constexpr int64_t kNoAllocInts = (int64_t)LARA_NOALLOC_SCRATCH_BYTES / 4;
static __device__ __align__(256) int32_t g_noAllocScratch[kNoAllocInts];

// header of the array, in ints (every part starts on a 256-byte boundary)
constexpr int64_t kNaDesc = 0;                          // PairDesc
constexpr int64_t kNaState = 128;                       // PairState
constexpr int64_t kNaSeqOff = 192;                      // int64_t[3]
constexpr int64_t kNaBody = 256;                        // sequences, bitmap, filter scratch
static_assert(sizeof(PairDesc) <= (kNaState - kNaDesc) * 4, "pair record outgrew its part of the scratch header");
static_assert(sizeof(PairState) <= (kNaSeqOff - kNaState) * 4, "solver state outgrew its part of the scratch header");

inline int64_t align64(int64_t x) { return (x + 63) & ~(int64_t)63; }

// device address of g_noAllocScratch per device: in host code the array's name is a host-side stand-in, only
// cudaGetSymbolAddress (with the device current) gives what a kernel can use
std::mutex g_noAllocMutex;
std::map<int, int32_t *> g_noAllocPtr;

int32_t * noalloc_scratch(int device)
{
    std::lock_guard<std::mutex> lock(g_noAllocMutex);
    auto it = g_noAllocPtr.find(device);
    if (it != g_noAllocPtr.end()) return it->second;
    CK(cudaSetDevice(device));
    int32_t * p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, g_noAllocScratch));
    if (!p) throw std::runtime_error("cudaGetSymbolAddress returned a null address for the scratch array");
    g_noAllocPtr[device] = p;
    return p;
}

// writes the header of the scratch array (everything that must not carry over from the call before) and copies the
// two sequences in
__global__ void noalloc_init_kernel(int32_t * __restrict__ scratch, PairDesc d, uint8_t const * __restrict__ seqA,
                                    uint8_t const * __restrict__ seqB)
{
    uint8_t * cat = (uint8_t *)(scratch + kNaBody);
    int const n = d.la + d.lb;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        cat[i] = i < d.la ? seqA[i] : seqB[i - d.la];
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        *(PairDesc *)(scratch + kNaDesc) = d;
        PairState st{};
        *(PairState *)(scratch + kNaState) = st;
        int64_t * off = (int64_t *)(scratch + kNaSeqOff);
        off[0] = 0; off[1] = d.la; off[2] = (int64_t)d.la + d.lb;
    }
}

// bitmap -> one byte per cell, optimum score -> the caller's word
__global__ void noalloc_expand_kernel(int32_t const * __restrict__ scratch, uint8_t * __restrict__ active,
                                      int32_t * __restrict__ optScore)
{
    PairDesc const & d = *(PairDesc const *)(scratch + kNaDesc);
    uint32_t const * bits = (uint32_t const *)(scratch + d.offBits);
    int64_t const cells = (int64_t)d.la * d.lb;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cells; i += (int64_t)gridDim.x * blockDim.x)
    {
        int const a = (int)(i / d.lb), b = (int)(i - (int64_t)a * d.lb);
        active[i] = (uint8_t)((bits[(size_t)a * d.bitWords + (b >> 5)] >> (b & 31)) & 1u);
    }
    if (optScore && blockIdx.x == 0 && threadIdx.x == 0)
        *optScore = ((PairState const *)(scratch + kNaState))->optScore;
}

} // namespace

extern "C" {

int lara_gpu_edge_filter_noalloc(lara_ctx * ctx, uint8_t const * d_seqA, int32_t lenA, uint8_t const * d_seqB, int32_t lenB,
                                 uint8_t * d_active, int32_t * d_opt_score)
{
    if (!ctx || !d_seqA || !d_seqB || lenA < 1 || lenB < 1 || !d_active) return set_err(ctx, LARA_E_INVALID, "bad argument");
    int64_t const bitWords = (lenB + 31) / 32;
    int64_t const seqInts = align64(((int64_t)lenA + lenB + 3) / 4);
    int64_t const bitInts = align64((int64_t)lenA * bitWords);
    int64_t const filterInts = (int64_t)lenA * lenB + 6ll * ((int64_t)lenA + 1);
    int64_t const need = kNaBody + seqInts + bitInts + filterInts;
    if (need > kNoAllocInts)
        return set_err(ctx, LARA_E_UNSUPPORTED,
                       "lara_gpu_edge_filter_noalloc: " + std::to_string(lenA) + " x " + std::to_string(lenB) + " nt needs " +
                       std::to_string(need * 4) + " bytes of scratch, the static array holds " +
                       std::to_string((long long)LARA_NOALLOC_SCRATCH_BYTES) + " (use lara_gpu_edge_filter)");
    Device & dv = ctx->devs[0];
    cudaStream_t const s = dv.stream;
    try
    {
        CK(cudaSetDevice(dv.id));
        int32_t * const scratch = noalloc_scratch(dv.id);
        PairDesc d{};
        d.la = lenA; d.lb = lenB; d.seqFirst = 0; d.seqSecond = 1; d.bitWords = (int32_t)bitWords;
        d.offBits = kNaBody + seqInts;
        d.offScratch = kNaBody + seqInts + bitInts;
        DevInputs in{};
        in.seqCat = (uint8_t const *)(scratch + kNaBody);
        in.seqOff = (int64_t const *)(scratch + kNaSeqOff);
        int const cpBlocks = (int)std::min<int64_t>(((int64_t)lenA + lenB + 255) / 256, 64);
        noalloc_init_kernel<<<cpBlocks, 256, 0, s>>>(scratch, d, d_seqA, d_seqB);
        CK(cudaGetLastError());
        CK(launch_edge_filter(1, (PairDesc *)(scratch + kNaDesc), (PairState *)(scratch + kNaState), scratch, in, ctx->dp,
                              lenB, s));
        int const exBlocks = (int)std::min<int64_t>(((int64_t)lenA * lenB + 255) / 256, 1184);
        noalloc_expand_kernel<<<exBlocks, 256, 0, s>>>(scratch, d_active, d_opt_score);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(s));
    }
    catch (std::exception const & ex)
    {
        return set_err(ctx, LARA_E_CUDA, ex.what());
    }
    return LARA_OK;
}

} // extern "C"
