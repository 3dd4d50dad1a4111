// trace_rt.h -- the tracer behind the B2_HOST_TRACE build of the kernel headers (test infrastructure).
//
// One host thread per CUDA thread (every lane independent), an emulated shared memory for all CTAs of a
// cluster, real blocking barriers, and happens-before tracked with vector clocks:
//   * bar.sync id, n / __syncthreads / __syncwarp / cluster.sync join the clocks of the threads they release;
//   * an mbarrier phase collects the clocks of everything that contributed to it -- the arrive.expect_tx and every
//     st.async that completed bytes on it -- and hands the join to each thread whose wait observes the phase;
//   * two accesses to the same 32-bit word from different threads, at least one a store, that are not ordered by
//     those edges are a race, whatever order the host scheduler ran them in.
// The including file defines TRACE_THREADS (threads per CTA x CTAs) and TRACE_CTA_THREADS before including.
#pragma once
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>
#include <vector>
#include <chrono>
#include <unistd.h>

thread_local TraceDim3 threadIdx{0, 0, 0}, blockIdx{0, 0, 0}, blockDim{1, 1, 1}, gridDim{1, 1, 1};

namespace trace {
constexpr int kThreads = TRACE_THREADS;
constexpr int kCtaThreads = TRACE_CTA_THREADS;
static std::vector<uint32_t> smem;                // emulated shared memory of all CTAs (words)
static size_t cta_words = 0;                      // words per CTA
static std::mutex mu;                             // one lock for the whole tracer
static std::condition_variable cv;
static int drop_site = 0;
static int deadlock_s = 20;                       // a wait longer than this is reported as a deadlock
static thread_local int tid = -1;                 // global thread id = cta * kCtaThreads + threadIdx.x
static inline int my_cta() { return tid / kCtaThreads; }

struct VC { uint32_t c[kThreads]; };
static VC vc[kThreads];
struct Access { int tid; uint32_t clk; };
struct Shadow { Access w{-1, 0}; std::vector<Access> r; };
static std::vector<Shadow> shadow;
static long races = 0, events = 0;
static std::map<std::pair<uint32_t, int>, int> race_words;

static inline bool ordered(const Access& a, int me) { return a.tid < 0 || a.tid == me || vc[me].c[a.tid] >= a.clk; }
static inline void report(uint32_t word, int kind)
{
    races++;
    if (race_words.size() < 8) race_words[{word, kind}]++;
}
static inline void join_into(VC& d, const VC& s) { for (int i = 0; i < kThreads; i++) d.c[i] = std::max(d.c[i], s.c[i]); }

// accesses (caller holds no lock)
static uint32_t ld_word(uint32_t w)
{
    std::lock_guard<std::mutex> lk(mu);
    if (w >= smem.size()) { printf("bad shared word %u\n", w); exit(2); }
    Shadow& s = shadow[w];
    events++;
    if (!ordered(s.w, tid)) report(w, 0);
    bool found = false;
    for (auto& a : s.r) if (a.tid == tid) { a.clk = vc[tid].c[tid]; found = true; }
    if (!found) s.r.push_back(Access{tid, vc[tid].c[tid]});
    return smem[w];
}
static void st_word_locked(uint32_t w, uint32_t v)
{
    if (w >= smem.size()) { printf("bad shared word %u\n", w); exit(2); }
    Shadow& s = shadow[w];
    events++;
    if (!ordered(s.w, tid)) report(w, 1);
    for (auto& a : s.r) if (!ordered(a, tid)) report(w, 2);
    s.r.clear();
    s.w = Access{tid, vc[tid].c[tid]};
    smem[w] = v;
}
static void st_word(uint32_t w, uint32_t v) { std::lock_guard<std::mutex> lk(mu); st_word_locked(w, v); }

// counted barriers: key identifies the barrier instance (CTA-local ids are offset by the CTA)
struct Bar { std::vector<int> arrived; uint64_t gen = 0; };
static std::map<long, Bar> bars;
static void barrier(long key, int n)
{
    std::unique_lock<std::mutex> lk(mu);
    Bar& b = bars[key];
    b.arrived.push_back(tid);
    if ((int)b.arrived.size() == n) {
        VC j{};
        for (int t : b.arrived) join_into(j, vc[t]);
        for (int t : b.arrived) { vc[t] = j; vc[t].c[t]++; }
        b.arrived.clear();
        b.gen++;
        cv.notify_all();
    } else {
        const uint64_t gen = b.gen;
        if (!cv.wait_for(lk, std::chrono::seconds(deadlock_s), [&] { return b.gen != gen; })) {
            printf("DEADLOCK: thread %d waits at barrier %ld (%zu of %d arrived); races so far %ld\n", tid, key, b.arrived.size(), n, races);
            fflush(stdout);
            _exit(3);
        }
    }
}

// mbarrier with transaction count (mbarrier.init / arrive.expect_tx / complete_tx by st.async / try_wait.parity)
struct MBar { uint32_t init = 0; int pending = 0; long tx = 0; uint32_t phase = 0; VC acc{}; VC done[2]{}; };
static std::map<uint32_t, MBar> mbars;            // key: byte address in the emulated shared memory
static void mbar_check_complete(MBar& m)
{
    if (m.pending == 0 && m.tx == 0) {
        m.done[m.phase & 1] = m.acc;
        m.acc = VC{};
        m.pending = (int)m.init;
        m.phase++;
        cv.notify_all();
    }
}
static void mbar_contribute(MBar& m)              // my clock flows into the phase; later events of mine are not covered
{
    join_into(m.acc, vc[tid]);
    vc[tid].c[tid]++;
}
}  // namespace trace

// ---- the hooks the B2_HOST_TRACE build of the kernel headers calls -------------------------------------------
namespace b2 {
uint32_t trace_smem_u32(const void* p) { return (uint32_t)((const char*)p - (const char*)trace::smem.data()); }
uint32_t* trace_cta_smem() { return trace::smem.data() + (size_t)trace::my_cta() * trace::cta_words; }
bool trace_sync_kept(int site) { return site != trace::drop_site; }
uint32_t trace_ld(uint32_t addr, int word)
{
    if (addr % 4) { printf("misaligned shared address %u\n", addr); exit(2); }
    return trace::ld_word(addr / 4 + word);
}
void trace_st(uint32_t addr, int word, uint32_t v)
{
    if (addr % 4) { printf("misaligned shared address %u\n", addr); exit(2); }
    trace::st_word(addr / 4 + word, v);
}
void trace_bar(int id, int nthreads) { trace::barrier((long)trace::my_cta() * 4096 + id, nthreads); }
void trace_warpsync() { trace::barrier(1000000 + trace::tid / 32, 32); }
int trace_cluster_rank() { return trace::my_cta(); }
void trace_cluster_sync() { trace::barrier(2000000, trace::kThreads); }
uint32_t trace_mapa(uint32_t addr, int rank)
{
    const uint32_t cta_bytes = (uint32_t)(trace::cta_words * 4);
    return addr % cta_bytes + (uint32_t)rank * cta_bytes;
}
void trace_mbar_init(uint32_t bar, uint32_t count)
{
    std::lock_guard<std::mutex> lk(trace::mu);
    trace::MBar& m = trace::mbars[bar];
    m.init = count; m.pending = (int)count; m.tx = 0; m.phase = 0;
}
void trace_mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    std::lock_guard<std::mutex> lk(trace::mu);
    trace::MBar& m = trace::mbars[bar];
    if (m.pending <= 0) { trace::report(bar / 4, 3); return; }          // arrival on a phase that has no arrival left: undefined
    m.tx += bytes;
    m.pending--;
    trace::mbar_contribute(m);
    trace::mbar_check_complete(m);
}
// st.async.v4 ... mbarrier::complete_tx::bytes: four remote words, then 16 bytes complete on the remote mbarrier
void trace_st_async(uint32_t remote_addr, int32_t a, int32_t b, int32_t c, int32_t d, uint32_t remote_bar)
{
    std::lock_guard<std::mutex> lk(trace::mu);
    const uint32_t w = remote_addr / 4;
    trace::st_word_locked(w, (uint32_t)a); trace::st_word_locked(w + 1, (uint32_t)b);
    trace::st_word_locked(w + 2, (uint32_t)c); trace::st_word_locked(w + 3, (uint32_t)d);
    trace::MBar& m = trace::mbars[remote_bar];
    m.tx -= 16;
    trace::mbar_contribute(m);
    trace::mbar_check_complete(m);
}
// cp.async.bulk global -> shared with mbarrier::complete_tx::bytes: the words are written on behalf of the issuing
// thread (the async proxy starts after the issue), then the bytes complete on the mbarrier
void trace_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    std::lock_guard<std::mutex> lk(trace::mu);
    const uint32_t* s = (const uint32_t*)src;
    for (uint32_t i = 0; i < bytes / 4; i++) trace::st_word_locked(dst / 4 + i, s[i]);
    trace::MBar& m = trace::mbars[bar];
    m.tx -= bytes;
    trace::mbar_contribute(m);
    trace::mbar_check_complete(m);
}
// try_wait.parity P: returns once the phase with parity P has completed
void trace_mbar_wait(uint32_t bar, uint32_t parity)
{
    std::unique_lock<std::mutex> lk(trace::mu);
    trace::MBar& m = trace::mbars[bar];
    if (!trace::cv.wait_for(lk, std::chrono::seconds(trace::deadlock_s), [&] { return (m.phase & 1u) != (parity & 1u); })) {
        printf("DEADLOCK: thread %d waits for parity %u of the mbarrier at %u (phase %u, %d arrivals and %ld bytes pending); races so far %ld\n",
               trace::tid, parity, bar, m.phase, m.pending, m.tx, trace::races);
        fflush(stdout);
        _exit(3);
    }
    trace::join_into(trace::vc[trace::tid], m.done[parity & 1u]);
}
}  // namespace b2

namespace trace {
// run `ctas` CTAs of `kCtaThreads` threads concurrently (one cluster); fn(cta, thread) is the kernel body
template <class F> static void launch(int ctas, F fn, int nthreads = kCtaThreads)
{
    std::vector<std::thread> th;
    for (int c = 0; c < ctas; c++)
        for (int t = 0; t < nthreads; t++)
            th.emplace_back([=] {
                tid = c * kCtaThreads + t;
                threadIdx = TraceDim3{(unsigned)t, 0, 0};
                blockIdx = TraceDim3{(unsigned)c, 0, 0};
                blockDim = TraceDim3{(unsigned)nthreads, 1, 1};
                gridDim = TraceDim3{(unsigned)ctas, 1, 1};
                fn(c, t);
            });
    for (auto& x : th) x.join();
}
static void print_races()
{
    for (auto& kv : race_words)
        printf("  e.g. word %u (%s)\n", kv.first.first,
               kv.first.second == 0 ? "read after unordered write" : kv.first.second == 1 ? "write after unordered write"
               : kv.first.second == 2 ? "write after unordered read" : "mbarrier arrival on a phase with no arrival left");
}
}  // namespace trace
