// b2fhe_defer.inl -- scattered-pointer batches and the deferred per-gate queue (included by
// b2fhe_api.cu after b2fhe_gates.inl).
//
// The reference's API is one gate per call on many streams (src/cufhe_gates_gpu.cu:148-665,
// test/test_gate_gpu.cc:68-83): one kernel launch of one block per gate.  Run that way this
// engine's batch-major kernels are starved (measured 1.9 k gates/s, tools/probe_streams.py).
// Instead the per-gate calls of the C++ shim are QUEUED here and executed at
// Synchronize()/StreamQuery()/explicit flush as a few batched launches, levelised by their
// data dependencies (a gate that consumes another queued gate's output runs one level
// later), which is legal because the reference only guarantees results after
// Synchronize() / a successful StreamQuery (include/cufhe_gpu.cuh:68-74).
#include <algorithm>
#include <unordered_map>
#include <unordered_set>

namespace {

template <typename T>
__global__ void k_gather(T* __restrict__ dst, const void* const* __restrict__ src, int count, int words)
{
    for (int g = blockIdx.x; g < count; g += gridDim.x) {
        const T* s = static_cast<const T*>(src[g]);
        if (!s) continue;
        for (int i = threadIdx.x; i < words; i += blockDim.x) dst[(size_t)g * words + i] = s[i];
    }
}
template <typename T>
__global__ void k_scatter(void* const* __restrict__ dst, const T* __restrict__ src, int count, int words)
{
    for (int g = blockIdx.x; g < count; g += gridDim.x) {
        T* d = static_cast<T*>(dst[g]);
        for (int i = threadIdx.x; i < words; i += blockDim.x) d[i] = src[(size_t)g * words + i];
    }
}

// host<->device copies of many small buffers as ONE transfer through a staging area:
// item = (device buffer, offset into the staging area, length), all in 16-bit units
struct CopyItem { void* dev; uint32_t off, n; };
template <bool TO_STAGE>
__global__ void k_stage_copy(uint16_t* __restrict__ stage, const CopyItem* __restrict__ items, int count)
{
    for (int g = blockIdx.x; g < count; g += gridDim.x) {
        const CopyItem it = items[g];
        uint16_t* d = static_cast<uint16_t*>(it.dev);
        uint16_t* s = stage + it.off;
        for (uint32_t i = threadIdx.x; i < it.n; i += blockDim.x)
            if (TO_STAGE) s[i] = d[i]; else d[i] = s[i];
    }
}

// gates over scattered ciphertext buffers: gather -> the batched path -> scatter
template <typename T>
int gate_ptrs(const DevCtx& c, cudaStream_t st, const uint8_t* ops, int level, void* const* out, const void* const* in0,
              const void* const* in1, const void* const* in2, int count)
{
    const int words = level == 0 ? kLvl0Words : kLvl1Words;
    Scratch S(st);
    std::vector<const void*> tab(4 * (size_t)count);
    for (int g = 0; g < count; g++) {
        REQUIRE(out[g] && in0[g], B2FHE_E_BAD_ARGUMENT);
        // two-input gates need in1, Mux / NMux also in2 (the rule of b2fhe_defer_gate): a null entry would make
        // the gather skip the slot and the blind rotation read uninitialised staging memory
        const int op = ops[g];
        REQUIRE(op < B2FHE_NUM_OPS, B2FHE_E_BAD_ARGUMENT);
        REQUIRE(op >= B2FHE_NOT || (in1 && in1[g]), B2FHE_E_BAD_ARGUMENT);
        REQUIRE((op != B2FHE_MUX && op != B2FHE_NMUX) || (in2 && in2[g]), B2FHE_E_BAD_ARGUMENT);
        tab[g] = in0[g];
        tab[count + g] = in1 ? in1[g] : nullptr;
        tab[2 * (size_t)count + g] = in2 ? in2[g] : nullptr;
        tab[3 * (size_t)count + g] = out[g];
    }
    const void** d_tab;
    if (int r = upload(S, &d_tab, tab)) return r;
    T* stage;
    if (int r = S.get((void**)&stage, 4 * (size_t)count * words * sizeof(T))) return r;
    const size_t n = (size_t)count * words;
    const int grid = count < 4096 ? count : 4096;
    for (int k = 0; k < 3; k++) {
        k_gather<T><<<grid, 128, 0, st>>>(stage + k * n, d_tab + (size_t)k * count, count, words);
        LAUNCHED();
    }
    if (int r = mixed_gate(c, st, ops, level, stage + 3 * n, stage, stage + n, stage + 2 * n, count)) return r;
    k_scatter<T><<<grid, 128, 0, st>>>((void* const*)(d_tab + 3 * (size_t)count), stage + 3 * n, count, words);
    LAUNCHED();
    return 0;
}

// ---- deferred queue ----------------------------------------------------------------------
struct Pending { uint8_t op, level; int lvl; void* out; const void* in[3]; };
struct PendingCopy { void* host; const void* dev; size_t bytes; int lvl = 0; };   // d2h: taken before gate level `lvl` runs
struct Pinned { void* p = nullptr; size_t cap = 0; std::atomic<int> busy{0}; };
// runs on the flush stream (cudaLaunchHostFunc) after the staged copy-back has landed
struct HostJob {
    struct Out { void* host; size_t off, bytes; };
    std::vector<Out> outs;
    Pinned* stage_in = nullptr;
    Pinned* stage_out = nullptr;
};
void CUDART_CB run_host_job(void* arg)
{
    HostJob* j = static_cast<HostJob*>(arg);
    for (const HostJob::Out& o : j->outs) memcpy(o.host, static_cast<const char*>(j->stage_out->p) + o.off, o.bytes);
    if (j->stage_in) j->stage_in->busy.store(0, std::memory_order_release);
    if (j->stage_out) j->stage_out->busy.store(0, std::memory_order_release);
    delete j;
}
struct DeferCtx {
    std::vector<Pending> q;
    std::vector<PendingCopy> d2h;
    std::vector<PendingCopy> h2d;                            // .host = source, .dev = destination
    std::unordered_map<const void*, size_t> h2d_index;       // destination -> entry of h2d
    std::unordered_map<const void*, size_t> d2h_index;       // host destination -> entry of d2h
    std::vector<Pinned*> pinned;                             // staging areas, reused
    std::unordered_set<const void*> produced;                // outputs of flushed gates, until the next completed wait
    std::unordered_set<const void*> backed;                  // flushed outputs whose copy-back into the caller's host ciphertext went with them
    std::unordered_map<const void*, int> wlevel, rlevel;     // buffer -> level of its pending writer / last reader
    std::unordered_map<cudaStream_t, cudaEvent_t> seen;      // user streams with work this queue depends on
    std::vector<cudaEvent_t> pool;                           // events of flushed entries, reused
    int maxlvl = 0;                                          // deepest dependency level queued
    cudaStream_t fs = nullptr;                               // flush stream
    cudaEvent_t done = nullptr;
};
DeferCtx g_defer[kMaxDev];
constexpr int kDeferAutoFlushPerSm = 32;
std::mutex g_defer_mu;

int acquire_pinned(DeferCtx& D, size_t bytes, Pinned*& out)
{
    for (Pinned* p : D.pinned)
        if (p->cap >= bytes && !p->busy.load(std::memory_order_acquire)) { p->busy.store(1); out = p; return 0; }
    Pinned* p = new Pinned;
    p->cap = bytes < (1u << 20) ? (1u << 20) : bytes + bytes / 2;
    CK(cudaMallocHost(&p->p, p->cap));
    p->busy.store(1);
    D.pinned.push_back(p);
    out = p;
    return 0;
}

void defer_reset(DeferCtx& D)
{
    D.q.clear(); D.d2h.clear(); D.h2d.clear(); D.h2d_index.clear(); D.d2h_index.clear(); D.wlevel.clear(); D.rlevel.clear();
    D.maxlvl = 0;
}

int defer_flush_body(int device, HostJob* job, bool& handed_over)
{
    DeferCtx& D = g_defer[device];
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (!D.fs) {
        CK(cudaStreamCreateWithFlags(&D.fs, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&D.done, cudaEventDisableTiming));
    }
    for (auto& kv : D.seen) CK(cudaStreamWaitEvent(D.fs, kv.second, 0));
    Scratch S(D.fs);
    std::vector<CopyItem> items;
    // ---- operands: host buffers -> one pinned staging area -> one transfer -> their device buffers
    if (!D.h2d.empty()) {
        size_t total = 0;
        items.reserve(D.h2d.size());
        for (const PendingCopy& pc : D.h2d) {
            items.push_back(CopyItem{const_cast<void*>(pc.dev), (uint32_t)(total / 2), (uint32_t)(pc.bytes / 2)});
            total += (pc.bytes + 15) & ~size_t(15);
        }
        if (int r = acquire_pinned(D, total, job->stage_in)) return r;
        char* hp = static_cast<char*>(job->stage_in->p);
        for (size_t i = 0; i < D.h2d.size(); i++) memcpy(hp + 2 * (size_t)items[i].off, D.h2d[i].host, D.h2d[i].bytes);
        uint16_t* d_stage;
        CopyItem* d_items;
        if (int r = S.get((void**)&d_stage, total)) return r;
        if (int r = upload(S, &d_items, items)) return r;
        CK(cudaMemcpyAsync(d_stage, hp, total, cudaMemcpyHostToDevice, D.fs));
        const int n = (int)items.size();
        k_stage_copy<false><<<n < 4096 ? n : 4096, 128, 0, D.fs>>>(d_stage, d_items, n);
        LAUNCHED();
    }
    // ---- results are snapshots: a copy-back queued after its producer (level L) is gathered into the
    // staging area before level L + 1 runs, so a later in-place update of the same buffer does not
    // leak into it (stream order of the reference); one transfer for all of them at the end
    uint16_t* d_out_stage = nullptr;
    CopyItem* d_out_items = nullptr;
    std::vector<int> first_of_lvl;          // items sorted by lvl: [first_of_lvl[l], first_of_lvl[l+1])
    size_t out_total = 0;
    int maxl = 0;
    for (const Pending& p : D.q) maxl = p.lvl > maxl ? p.lvl : maxl;
    if (!D.d2h.empty()) {
        std::vector<size_t> order(D.d2h.size());
        for (size_t i = 0; i < order.size(); i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](size_t x, size_t y) { return D.d2h[x].lvl < D.d2h[y].lvl; });
        items.clear();
        items.reserve(order.size());
        first_of_lvl.assign(maxl + 3, 0);
        for (size_t i : order) {
            const PendingCopy& pc = D.d2h[i];
            const int l = pc.lvl > maxl + 1 ? maxl + 1 : pc.lvl;
            first_of_lvl[l + 1]++;
            items.push_back(CopyItem{const_cast<void*>(pc.dev), (uint32_t)(out_total / 2), (uint32_t)(pc.bytes / 2)});
            job->outs.push_back(HostJob::Out{pc.host, out_total, pc.bytes});
            out_total += (pc.bytes + 15) & ~size_t(15);
        }
        for (int l = 0; l <= maxl + 1; l++) first_of_lvl[l + 1] += first_of_lvl[l];
        if (int r = acquire_pinned(D, out_total, job->stage_out)) return r;
        if (int r = S.get((void**)&d_out_stage, out_total)) return r;
        if (int r = upload(S, &d_out_items, items)) return r;
    }
    auto gather_results = [&](int l) -> int {
        if (first_of_lvl.empty()) return 0;
        const int lo = first_of_lvl[l], n = first_of_lvl[l + 1] - lo;
        if (n <= 0) return 0;
        k_stage_copy<true><<<n < 4096 ? n : 4096, 128, 0, D.fs>>>(d_out_stage, d_out_items + lo, n);
        LAUNCHED();
        return 0;
    };
    // ---- the gates, level by level
    std::vector<uint8_t> ops;
    std::vector<void*> out;
    std::vector<const void*> in[3];
    for (int l = 0; l <= maxl; l++) {
        if (int r = gather_results(l)) return r;
        for (int level = 0; level < 2; level++) {
            ops.clear(); out.clear(); in[0].clear(); in[1].clear(); in[2].clear();
            for (const Pending& p : D.q)
                if (p.lvl == l && p.level == level) {
                    ops.push_back(p.op); out.push_back(p.out);
                    for (int k = 0; k < 3; k++) in[k].push_back(p.in[k]);
                }
            if (ops.empty()) continue;
            const int n = (int)ops.size();
            const int r = level == 0 ? gate_ptrs<uint16_t>(*c, D.fs, ops.data(), 0, out.data(), in[0].data(), in[1].data(), in[2].data(), n)
                                     : gate_ptrs<uint32_t>(*c, D.fs, ops.data(), 1, out.data(), in[0].data(), in[1].data(), in[2].data(), n);
            if (r) return r;
        }
    }
    if (int r = gather_results(maxl + 1)) return r;
    if (out_total) CK(cudaMemcpyAsync(job->stage_out->p, d_out_stage, out_total, cudaMemcpyDeviceToHost, D.fs));
    CK(cudaLaunchHostFunc(D.fs, run_host_job, job));
    handed_over = true;
    CK(cudaEventRecord(D.done, D.fs));
    // later work on the user streams is ordered after this flush ("same stream = issue order")
    for (auto& kv : D.seen) {
        // a stream the caller has already destroyed (legal with work pending) has nothing left to order
        if (cudaStreamWaitEvent(kv.first, D.done, 0) != cudaSuccess) (void)cudaGetLastError();
        D.pool.push_back(kv.second);
    }
    // the user may destroy its streams after Synchronize(): keep no handle past the flush
    D.seen.clear();
    for (const Pending& p : D.q) D.produced.insert(p.out);
    for (const PendingCopy& pc : D.d2h) {          // copy-backs taken right after the buffer's (last) producer
        const auto w = D.wlevel.find(pc.dev);
        if (w == D.wlevel.end() || pc.lvl == w->second + 1) D.backed.insert(pc.dev);
    }
    return 0;
}

int defer_flush(int device)
{
    DeferCtx& D = g_defer[device];
    if (D.q.empty() && D.d2h.empty() && D.h2d.empty()) return 0;
    HostJob* job = new HostJob;
    bool handed_over = false;
    const int r = defer_flush_body(device, job, handed_over);
    if (!handed_over) {       // failed before the host function took ownership
        if (job->stage_in) job->stage_in->busy.store(0);
        if (job->stage_out) job->stage_out->busy.store(0);
        delete job;
    }
    defer_reset(D);           // on failure too: a broken queue must not be retried for ever
    return r;
}

// after a wait that covered every flushed gate, host copies are authoritative again
void defer_mark_idle(int device) { g_defer[device].produced.clear(); g_defer[device].backed.clear(); }
void defer_mark_idle_all()
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    for (int d = 0; d < kMaxDev; d++) defer_mark_idle(d);
}
// the caller may have waited through raw CUDA (cudaDeviceSynchronize / cudaStreamSynchronize, as the
// reference's test/test_intensive.cc:43 does): once the last flush has completed, host copies are
// authoritative again whoever did the waiting
void defer_refresh_idle(DeferCtx& D, int device)
{
    if (D.produced.empty() || !D.done) return;
    if (cudaEventQuery(D.done) == cudaSuccess) defer_mark_idle(device);
    else (void)cudaGetLastError();
}

int defer_flush_all()
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    for (int d = 0; d < kMaxDev; d++)
        if (g_ctx[d].ready)
            if (int r = defer_flush(d)) return r;
    return 0;
}

int defer_destroy_all()
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    for (int d = 0; d < kMaxDev; d++) {
        DeferCtx& D = g_defer[d];
        if (!D.fs && D.seen.empty() && D.pool.empty() && D.pinned.empty()) continue;
        cudaSetDevice(d);
        for (auto& kv : D.seen) cudaEventDestroy(kv.second);
        for (cudaEvent_t e : D.pool) cudaEventDestroy(e);
        for (Pinned* p : D.pinned) { cudaFreeHost(p->p); delete p; }
        if (D.done) cudaEventDestroy(D.done);
        if (D.fs) cudaStreamDestroy(D.fs);
        D = DeferCtx();
    }
    return 0;
}

}  // namespace

extern "C" {

int b2fhe_gate_batch_ptrs(int device, b2fhe_stream_t stream, const uint8_t* ops_host, int level, void* const* out,
                          const void* const* in0, const void* const* in1, const void* const* in2, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE((level == 0 || level == 1) && count < (1u << 28), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(ops_host && out && in0, B2FHE_E_BAD_ARGUMENT);
    return level == 0 ? gate_ptrs<uint16_t>(*c, (cudaStream_t)stream, ops_host, 0, out, in0, in1, in2, (int)count)
                      : gate_ptrs<uint32_t>(*c, (cudaStream_t)stream, ops_host, 1, out, in0, in1, in2, (int)count);
}

int b2fhe_defer_gate(int device, b2fhe_stream_t user_stream, int op, int level, void* out, const void* in0,
                     const void* in1, const void* in2)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(op >= 0 && op < B2FHE_NUM_OPS && (level == 0 || level == 1) && out && in0, B2FHE_E_BAD_ARGUMENT);
    const bool two = op < B2FHE_MUX, three = op == B2FHE_MUX || op == B2FHE_NMUX;
    REQUIRE((!two && !three) || in1, B2FHE_E_BAD_ARGUMENT);
    REQUIRE(!three || in2, B2FHE_E_BAD_ARGUMENT);
    DeferCtx& D = g_defer[device];
    defer_refresh_idle(D, device);
    // a long run of INDEPENDENT gates (the reference's benchmark loop) starts executing while the
    // caller is still issuing: two full waves of the blind-rotation kernel lose nothing by going
    // now.  Checked BEFORE queueing, so that the copy-back of every flushed gate goes with it.
    if (D.maxlvl == 0 && (int)D.q.size() >= kDeferAutoFlushPerSm * c->sm_count)
        if (int r = defer_flush(device)) return r;
    Pending p{};
    p.op = (uint8_t)op; p.level = (uint8_t)level; p.out = out;
    p.in[0] = in0; p.in[1] = (two || three) ? in1 : nullptr; p.in[2] = three ? in2 : nullptr;
    int lvl = 0;
    for (int k = 0; k < 3; k++)
        if (p.in[k]) { auto it = D.wlevel.find(p.in[k]); if (it != D.wlevel.end() && it->second + 1 > lvl) lvl = it->second + 1; }
    { auto it = D.wlevel.find(out); if (it != D.wlevel.end() && it->second + 1 > lvl) lvl = it->second + 1; }   // write after write
    { auto it = D.rlevel.find(out); if (it != D.rlevel.end() && it->second > lvl) lvl = it->second; }           // write after read: same level is safe (inputs are gathered first)
    p.lvl = lvl;
    D.wlevel[out] = lvl;
    for (int k = 0; k < 3; k++)
        if (p.in[k]) { int& r = D.rlevel[p.in[k]]; if (lvl > r) r = lvl; }
    D.q.push_back(p);
    // whatever the caller queued on its stream before this call (H2D of the operands) must be visible
    cudaStream_t us = (cudaStream_t)user_stream;
    auto it = D.seen.find(us);
    if (it == D.seen.end()) {
        cudaEvent_t ev;
        if (!D.pool.empty()) { ev = D.pool.back(); D.pool.pop_back(); }
        else CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        it = D.seen.emplace(us, ev).first;
    }
    CK(cudaEventRecord(it->second, us));
    if (lvl > D.maxlvl) D.maxlvl = lvl;
    return 0;
}

int b2fhe_defer_d2h(int device, void* host_dst, const void* dev_src, size_t bytes)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev && host_dst && dev_src && bytes % 2 == 0 && bytes < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    DeferCtx& D = g_defer[device];
    // a repeated host destination needs only the last copy
    const auto w = D.wlevel.find(dev_src);
    const int lvl = w == D.wlevel.end() ? 0 : w->second + 1;        // right after its queued producer
    auto it = D.d2h_index.find(host_dst);
    if (it != D.d2h_index.end()) { D.d2h[it->second] = PendingCopy{host_dst, dev_src, bytes, lvl}; return 0; }
    D.d2h_index.emplace(host_dst, D.d2h.size());
    D.d2h.push_back(PendingCopy{host_dst, dev_src, bytes, lvl});
    return 0;
}

/* queued host-to-device copy of a gate operand, performed when the queue runs (all of them as
 * one transfer).  When a queued or still-running gate writes `dev_dst`:
 *   - if a copy-back of that buffer into the caller's host ciphertext is queued after that gate (or went
 *     with it), host and device will agree: the upload is dropped;
 *   - otherwise (the producer was a non-copying g* gate) the reference's stream order makes the HOST bytes
 *     win: the queue is flushed and waited for, then the upload is queued like any other. */
int b2fhe_defer_h2d(int device, void* dev_dst, const void* host_src, size_t bytes)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev && host_src && dev_dst && bytes % 2 == 0 && bytes < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    DeferCtx& D = g_defer[device];
    defer_refresh_idle(D, device);
    const auto w = D.wlevel.find(dev_dst);
    if (w != D.wlevel.end() || D.produced.count(dev_dst)) {
        bool agrees = w == D.wlevel.end() && D.backed.count(dev_dst) != 0;
        if (w != D.wlevel.end())
            for (const PendingCopy& pc : D.d2h)
                if (pc.dev == dev_dst && pc.lvl == w->second + 1) { agrees = true; break; }
        if (agrees) return 0;
        if (int r = defer_flush(device)) return r;
        if (D.done) { CK(cudaSetDevice(device)); CK(cudaEventSynchronize(D.done)); }
        defer_mark_idle(device);
    }
    auto it = D.h2d_index.find(dev_dst);
    if (it != D.h2d_index.end()) { D.h2d[it->second] = PendingCopy{const_cast<void*>(host_src), dev_dst, bytes}; return 0; }
    D.h2d_index.emplace(dev_dst, D.h2d.size());
    D.h2d.push_back(PendingCopy{const_cast<void*>(host_src), dev_dst, bytes});
    return 0;
}

int b2fhe_defer_flush(int device)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev, B2FHE_E_BAD_ARGUMENT);
    return defer_flush(device);
}

/* flush, then make `stream` wait for everything flushed so far (orders immediate calls and
 * explicit copies after the deferred gates) */
int b2fhe_defer_fence(int device, b2fhe_stream_t stream)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev, B2FHE_E_BAD_ARGUMENT);
    if (int r = defer_flush(device)) return r;
    DeferCtx& D = g_defer[device];
    if (D.done) { CK(cudaSetDevice(device)); CK(cudaStreamWaitEvent((cudaStream_t)stream, D.done, 0)); }
    return 0;
}

/* the fence, but only when a queued gate reads or writes `buffer` (an explicit copy into a
 * ciphertext must not overtake a queued gate that still uses it) */
int b2fhe_defer_fence_buffer(int device, b2fhe_stream_t stream, const void* buffer)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev, B2FHE_E_BAD_ARGUMENT);
    DeferCtx& D = g_defer[device];
    if (D.wlevel.find(buffer) == D.wlevel.end() && D.rlevel.find(buffer) == D.rlevel.end()) return 0;
    if (int r = defer_flush(device)) return r;
    CK(cudaSetDevice(device));
    CK(cudaStreamWaitEvent((cudaStream_t)stream, D.done, 0));
    return 0;
}

int b2fhe_defer_pending(int device)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    if (device < 0 || device >= kMaxDev) return 0;
    DeferCtx& D = g_defer[device];
    int inflight = 0;       // a flush that is still running may yet write a caller's host / device buffer (~Ctxt asks)
    if (D.done) {
        if (cudaEventQuery(D.done) == cudaErrorNotReady) inflight = 1;
        (void)cudaGetLastError();
    }
    return (int)(D.q.size() + D.d2h.size() + D.h2d.size()) + inflight;
}

/* flush, then 0 = everything flushed so far has completed, 1 = still running */
int b2fhe_defer_query(int device, int* busy)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev && busy, B2FHE_E_BAD_ARGUMENT);
    if (int r = defer_flush(device)) return r;
    DeferCtx& D = g_defer[device];
    *busy = 0;
    if (D.fs) {
        CK(cudaSetDevice(device));
        const cudaError_t e = cudaStreamQuery(D.fs);
        if (e == cudaErrorNotReady) *busy = 1;
        else if (e != cudaSuccess) return fail((int)e, "cudaStreamQuery", __FILE__, __LINE__);
    }
    if (!*busy) defer_mark_idle(device);
    return 0;
}

/* StreamQuery for ONE user stream: flushes only when `stream` itself has queued work (a polling loop over
 * many streams -- test/test_intensive.cc:21-57 -- then keeps forming full batches instead of one-gate
 * flushes); busy = work flushed so far is still running. */
int b2fhe_defer_query_stream(int device, b2fhe_stream_t stream, int* busy)
{
    std::lock_guard<std::mutex> lk(g_defer_mu);
    REQUIRE(device >= 0 && device < kMaxDev && busy, B2FHE_E_BAD_ARGUMENT);
    DeferCtx& D = g_defer[device];
    // (copies queued without any gate carry no stream: they go with the first query)
    if (D.seen.find((cudaStream_t)stream) != D.seen.end() || (D.q.empty() && (!D.d2h.empty() || !D.h2d.empty())))
        if (int r = defer_flush(device)) return r;
    *busy = 0;
    if (D.fs) {
        CK(cudaSetDevice(device));
        const cudaError_t e = cudaStreamQuery(D.fs);
        if (e == cudaErrorNotReady) *busy = 1;
        else if (e != cudaSuccess) return fail((int)e, "cudaStreamQuery", __FILE__, __LINE__);
    }
    if (!*busy && D.q.empty()) defer_mark_idle(device);
    return 0;
}

}  // extern "C"
