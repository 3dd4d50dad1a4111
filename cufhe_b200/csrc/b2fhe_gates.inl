// b2fhe_gates.inl -- gate / stage entry points of the C ABI (included by b2fhe_api.cu).
// Host-side composition that the reference does inside one kernel per gate
// (__HomGate__ both orders, Mux/NMux, Not/Copy: src/bootstrap_gpu.cu:383-780).

namespace {

template <int G> int launch_brs_g(const DevCtx& c, const BrParams& P, cudaStream_t st)
{
    const int grid = (P.count - P.first + G - 1) / G;
    k_blind_rotate_s<G><<<grid, 128 * G, brs_smem_bytes(G), st>>>(P, reinterpret_cast<const TeamSTables*>(c.tabs));
    LAUNCHED();
    return 0;
}
// Kernel choice is a function of the batch size and of explicit settings only (b2fhe_set_br_variant,
// b2fhe_set_latency_mode): the same call always runs the same kernels.
int launch_br(const DevCtx& c, BrParams P, cudaStream_t st)
{
    if (P.count <= 0) return 0;
    P.bk = c.bk;
    int g = g_gates_per_cta;
    const int per_sm = (P.count + c.sm_count - 1) / c.sm_count;
    const bool small_ok = g_br_variant == 0 && g == 0 && g_latency_mode;
    // up to 70 gates: one gate per cluster of 4 CTAs x 128 threads (70 clusters are resident at once, two CTAs
    // per SM); hand-overs by st.async + mbarrier
    if (g_br_variant == 5 || (small_ok && P.count <= g_br_cluster_max)) {
        k_blind_rotate_cl4s<false><<<P.count * kCl4Ctas, kCl4Threads, kCl4SmemBytes, st>>>(P, reinterpret_cast<const PrimeSTables*>(c.tabs));
        LAUNCHED();
        return 0;
    }
    // schedule fuzzing builds (random sleeps around every synchronisation; same bits): 6 = cluster kernel, 41 = throughput kernel
    if (g_br_variant == 6) {
        k_blind_rotate_cl4s<true><<<P.count * kCl4Ctas, kCl4Threads, kCl4SmemBytes, st>>>(P, reinterpret_cast<const PrimeSTables*>(c.tabs));
        LAUNCHED();
        return 0;
    }
    if (g_br_variant == 41) {
        if (g == 1) k_blind_rotate_s<1, false, ABL_JIT><<<P.count - P.first, 128, brs_smem_bytes(1), st>>>(P, reinterpret_cast<const TeamSTables*>(c.tabs));
        else k_blind_rotate_s<4, false, ABL_JIT><<<(P.count - P.first + 3) / 4, 512, brs_smem_bytes(4), st>>>(P, reinterpret_cast<const TeamSTables*>(c.tabs));
        LAUNCHED();
        return 0;
    }
    if (g_br_variant == 48) {      // key row 0 at every step (timing experiment: key traffic vs arithmetic; wrong results)
        const int grid = (P.count - P.first + 3) / 4;
        k_blind_rotate_s<4, true><<<grid, 512, brs_smem_bytes(4), st>>>(P, reinterpret_cast<const TeamSTables*>(c.tabs));
        LAUNCHED();
        return 0;
    }
    // the throughput kernel: four gates per CTA (16 warps at 128 registers fill the register file)
    if (g == 0 && g_br_tail_split) {
        // Waves of 4 gates per SM take the same time whether full or not; a last partial wave of at
        // most 3 gates per SM runs as its own launch with fewer gates per CTA.
        const int wave = 4 * c.sm_count;
        const int full = (P.count / wave) * wave, rest = P.count - full;
        if (full > 0 && rest > 0 && rest <= 3 * c.sm_count) {
            BrParams M = P;
            M.count = full;
            if (int r = launch_brs_g<4>(c, M, st)) return r;
            P.first = full;
            // 1 or 2 gates per SM: one-gate CTAs (two fit an SM)
            if (rest <= 2 * c.sm_count) return launch_brs_g<1>(c, P, st);
            return launch_brs_g<3>(c, P, st);
        }
    }
    if (g == 0) g = per_sm >= 4 ? 4 : per_sm == 3 ? 3 : 1;      // two one-gate CTAs share an SM faster than one two-gate CTA
    switch (g) {
        case 1: return launch_brs_g<1>(c, P, st);
        case 2: return launch_brs_g<2>(c, P, st);
        case 3: return launch_brs_g<3>(c, P, st);
        default: return launch_brs_g<4>(c, P, st);
    }
}
// work items of a uniform lvl0 Mux / NMux batch: [0, count) = inc AND in1, [count, 2 count) = (NOT inc) AND in0
__global__ void k_mux_items(BrItem* items, int count, int off)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * count) return;
    const int g = i < count ? i : i - count;
    BrItem b{};
    b.g0 = b.g1 = (uint32_t)g; b.ca = i < count ? 1 : -1; b.cb = 1; b.src1 = i < count ? 0 : 1; b.off = off; b.slot = (uint32_t)i;
    items[i] = b;
}
int launch_ks(const DevCtx& c, KsParams P, cudaStream_t st)
{
    if (P.count <= 0) return 0;
    P.T = c.ksk;
    if (g_ks_variant == 1 || (g_ks_variant == 0 && P.count <= g_ks_lat_max_per_sm * c.sm_count)) {
        // small batch: 16 warps per gate instead of one; 32 warps up to 4 gates per SM (half the chain of
        // L2 round trips: single NAND 1.51 -> 1.49 ms warm, 1.60 -> 1.53 ms from a flushed L2, profiles/r2_probe_lat_ks.log)
        if (P.count <= 4 * c.sm_count) k_keyswitch_lat<7, 32><<<P.count, 1024, 0, st>>>(P);
        else k_keyswitch_lat<7><<<P.count, 32 * kKsLatWarps, 0, st>>>(P);
        LAUNCHED();
        return 0;
    }
    if (g_ks_variant == 3) {      // A/B of the latency kernel: a branch per digit (10 key words per thread in flight), 16 warps
        k_keyswitch_lat<0><<<P.count, 32 * kKsLatWarps, 0, st>>>(P);
        LAUNCHED();
        return 0;
    }
    // in-step variants (A/B): 10 + (warps per CTA: 0 = 8, 1 = 16, 2 = 32) * 3 + (barrier every 1 / 2 / 4 coefficients)
    if (g_ks_variant == 0) {      // measured fastest with the 16x2 adds (profiles/r2_probe_ks_viadd.log): 16 gates per CTA, barrier every 2 coefficients
        k_keyswitch_step<16, 2><<<(P.count + 15) / 16, 32 * 16, 16 * kN * 2, st>>>(P);
        LAUNCHED();
        return 0;
    }
    if (g_ks_variant >= 10 && g_ks_variant < 22) {
#define KS_STEP(Wn, Sj) do { k_keyswitch_step<Wn, Sj><<<(P.count + Wn - 1) / Wn, 32 * Wn, Wn * kN * 2, st>>>(P); LAUNCHED(); return 0; } while (0)
        switch (g_ks_variant) {
            case 10: KS_STEP(8, 1);
            case 11: KS_STEP(8, 2);
            case 12: KS_STEP(8, 4);
            case 13: KS_STEP(16, 1);
            case 14: KS_STEP(16, 2);
            case 15: KS_STEP(16, 4);
            case 16: KS_STEP(32, 1);
            case 17: KS_STEP(32, 2);
            case 18: KS_STEP(32, 4);
            case 19: KS_STEP(32, 8);
            case 20: KS_STEP(32, 16);
            default: KS_STEP(16, 8);
        }
#undef KS_STEP
    }
    k_keyswitch<<<(P.count + kKsGates - 1) / kKsGates, 32 * kKsGates, 0, st>>>(P);
    LAUNCHED();
    return 0;
}
template <typename T>
int launch_affine(T* out, const T* A, const T* B, const AffItem* items, int count, int words, int ca, int cb,
                  uint32_t off, cudaStream_t st)
{
    if (count <= 0) return 0;
    k_affine<T><<<count < 65535 ? count : 65535, 128, 0, st>>>(out, A, B, items, count, words, ca, cb, off);
    LAUNCHED();
    return 0;
}

// stream-ordered scratch
struct Scratch {
    cudaStream_t st;
    std::vector<void*> ptrs;
    explicit Scratch(cudaStream_t s) : st(s) {}
    int get(void** p, size_t bytes)
    {
        CK(cudaMallocAsync(p, bytes ? bytes : 16, st));
        ptrs.push_back(*p);
        return 0;
    }
    ~Scratch() { for (void* p : ptrs) cudaFreeAsync(p, st); }
};

int uniform_gate(const DevCtx& c, cudaStream_t st, int op, int level, void* out, const void* in0, const void* in1,
                 const void* in2, int count)
{
    Scratch S(st);
    if (op == B2FHE_NOT || op == B2FHE_COPY) {
        const int ca = op == B2FHE_NOT ? -1 : 1;
        if (level == 0) return launch_affine<uint16_t>((uint16_t*)out, (const uint16_t*)in0, nullptr, nullptr, count, kLvl0Words, ca, 0, 0u, st);
        return launch_affine<uint32_t>((uint32_t*)out, (const uint32_t*)in0, nullptr, nullptr, count, kLvl1Words, ca, 0, 0u, st);
    }
    BrParams B{};
    KsParams K{};
    B.count = K.count = count;
    B.mu = kMu1;
    if (op < B2FHE_MUX) {
        int ca, cb, so;
        gate_coeffs(op, ca, cb, so);
        REQUIRE(in1 != nullptr, B2FHE_E_BAD_ARGUMENT);
        if (level == 0) {       // BR(pre-add) -> SE -> IKS   (src/bootstrap_gpu.cu:402-421)
            uint32_t* t1;
            if (int r = S.get((void**)&t1, (size_t)count * kLvl1Words * 4)) return r;
            B.in0 = (const uint16_t*)in0; B.in1 = (const uint16_t*)in1; B.out = t1;
            B.ca = ca; B.cb = cb; B.off = so * (int)kMu0;
            if (int r = launch_br(c, B, st)) return r;
            K.A = t1; K.out = (uint16_t*)out; K.ca = 1; K.cb = 0; K.off = 0u;
            return launch_ks(c, K, st);
        }
        // IKS(pre-add) -> BR -> SE   (src/bootstrap_gpu.cu:383-400)
        uint16_t* t0;
        if (int r = S.get((void**)&t0, (size_t)count * kLvl0Words * 2)) return r;
        K.A = (const uint32_t*)in0; K.B = (const uint32_t*)in1; K.out = t0;
        K.ca = ca; K.cb = cb; K.off = (uint32_t)so * kMu1;
        if (int r = launch_ks(c, K, st)) return r;
        B.in0 = t0; B.out = (uint32_t*)out; B.ca = 1; B.cb = 0; B.off = 0;
        return launch_br(c, B, st);
    }
    // MUX / NMUX: in0 = inc, in1 = in1, in2 = in0 of the reference signature
    REQUIRE(in1 != nullptr && in2 != nullptr, B2FHE_E_BAD_ARGUMENT);
    const int sg = op == B2FHE_NMUX ? -1 : 1;
    if (level == 0) {           // src/bootstrap_gpu.cu:515-588
        uint32_t* ta;
        if (int r = S.get((void**)&ta, 2 * (size_t)count * kLvl1Words * 4)) return r;
        // both branches as ONE blind-rotation launch of 2 * count work items (one wave quantisation, not two):
        // item g = inc AND in1 -> ta[g], item count + g = (NOT inc) AND in0 -> ta[count + g]
        BrItem* d_items;
        if (int r = S.get((void**)&d_items, 2 * (size_t)count * sizeof(BrItem))) return r;
        k_mux_items<<<(2 * count + 255) / 256, 256, 0, st>>>(d_items, count, -(int)kMu0);
        LAUNCHED();
        B.in0 = (const uint16_t*)in0; B.in1 = (const uint16_t*)in1; B.in2 = (const uint16_t*)in2; B.out = ta;
        B.items = d_items; B.count = 2 * count;
        if (int r = launch_br(c, B, st)) return r;
        K.A = ta; K.B = ta + (size_t)count * kLvl1Words; K.out = (uint16_t*)out; K.ca = sg; K.cb = sg; K.off = (uint32_t)sg * kMu1;
        return launch_ks(c, K, st);
    }
    // src/bootstrap_gpu.cu:706-780
    uint16_t *t0a, *t0b;
    uint32_t *ea, *eb;
    if (int r = S.get((void**)&t0a, (size_t)count * kLvl0Words * 2)) return r;
    if (int r = S.get((void**)&t0b, (size_t)count * kLvl0Words * 2)) return r;
    if (int r = S.get((void**)&ea, (size_t)count * kLvl1Words * 4)) return r;
    if (int r = S.get((void**)&eb, (size_t)count * kLvl1Words * 4)) return r;
    K.A = (const uint32_t*)in0; K.B = (const uint32_t*)in1; K.out = t0a; K.ca = 1; K.cb = 1; K.off = 0u - kMu1;
    if (int r = launch_ks(c, K, st)) return r;
    K.B = (const uint32_t*)in2; K.out = t0b; K.ca = -1;
    if (int r = launch_ks(c, K, st)) return r;
    B.in0 = t0a; B.out = ea; B.ca = 1; B.cb = 0; B.off = 0;
    if (int r = launch_br(c, B, st)) return r;
    B.in0 = t0b; B.out = eb;
    if (int r = launch_br(c, B, st)) return r;
    return launch_affine<uint32_t>((uint32_t*)out, ea, eb, nullptr, count, kLvl1Words, sg, sg, (uint32_t)sg * kMu1, st);
}

template <typename T> int upload(Scratch& S, T** dev, const std::vector<T>& host)
{
    if (int r = S.get((void**)dev, host.size() * sizeof(T))) return r;
    if (!host.empty()) CK(cudaMemcpyAsync(*dev, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice, S.st));
    return 0;
}

// per-gate op codes: build work-item lists on the host, then the same three kernels
int mixed_gate(const DevCtx& c, cudaStream_t st, const uint8_t* ops, int level, void* out, const void* in0,
               const void* in1, const void* in2, int count)
{
    Scratch S(st);
    std::vector<BrItem> br;
    std::vector<KsItem> ks;
    std::vector<AffItem> af;    // level 0: Not/Copy only; level 1: every gate's final write
    br.reserve(count + count / 4);
    ks.reserve(count + count / 4);
    for (int g = 0; g < count; g++) {
        const int op = ops[g];
        REQUIRE(op < B2FHE_NUM_OPS, B2FHE_E_BAD_ARGUMENT);
        if (op == B2FHE_NOT || op == B2FHE_COPY) {
            AffItem a{}; a.s0 = a.s1 = a.slot = (uint32_t)g; a.ca = op == B2FHE_NOT ? -1 : 1; a.cb = 0; a.off = 0;
            a.pad[0] = 1;   // marks "source is the input array" for level 1
            af.push_back(a);
            continue;
        }
        int ca, cb, so;
        const bool mux = op >= B2FHE_MUX;
        const int sg = op == B2FHE_NMUX ? -1 : 1;
        if (!mux) gate_coeffs(op, ca, cb, so);
        if (level == 0) {
            KsItem k{}; k.slot = (uint32_t)g;
            if (!mux) {
                BrItem b{}; b.g0 = b.g1 = (uint32_t)g; b.ca = (int8_t)ca; b.cb = (int8_t)cb; b.src1 = 0;
                b.off = so * (int)kMu0; b.slot = (uint32_t)br.size();
                k.s0 = k.s1 = b.slot; k.ca = 1; k.cb = 0; k.off = 0;
                br.push_back(b);
            } else {
                BrItem b{}; b.g0 = b.g1 = (uint32_t)g; b.ca = 1; b.cb = 1; b.src1 = 0; b.off = -(int)kMu0;
                b.slot = (uint32_t)br.size(); k.s0 = b.slot; br.push_back(b);
                b.ca = -1; b.src1 = 1; b.slot = (uint32_t)br.size(); k.s1 = b.slot; br.push_back(b);
                k.ca = (int8_t)sg; k.cb = (int8_t)sg; k.off = (uint32_t)sg * kMu1;
            }
            ks.push_back(k);
        } else {
            AffItem a{}; a.slot = (uint32_t)g;
            if (!mux) {
                KsItem k{}; k.s0 = k.s1 = (uint32_t)g; k.ca = (int8_t)ca; k.cb = (int8_t)cb; k.off = (uint32_t)so * kMu1;
                k.slot = (uint32_t)ks.size(); k.pad[0] = 0;
                a.s0 = a.s1 = k.slot; a.ca = 1; a.cb = 0; a.off = 0;
                ks.push_back(k);
            } else {
                KsItem k{}; k.s0 = k.s1 = (uint32_t)g; k.ca = 1; k.cb = 1; k.off = 0u - kMu1; k.pad[0] = 0;
                k.slot = (uint32_t)ks.size(); a.s0 = k.slot; ks.push_back(k);
                k.ca = -1; k.pad[0] = 1; k.slot = (uint32_t)ks.size(); a.s1 = k.slot; ks.push_back(k);
                a.ca = (int8_t)sg; a.cb = (int8_t)sg; a.off = (uint32_t)sg * kMu1;
            }
            af.push_back(a);
        }
    }
    BrParams B{};
    KsParams K{};
    B.mu = kMu1;
    if (level == 0) {
        BrItem* d_br; KsItem* d_ks; AffItem* d_af; uint32_t* t1;
        if (int r = upload(S, &d_br, br)) return r;
        if (int r = upload(S, &d_ks, ks)) return r;
        if (int r = upload(S, &d_af, af)) return r;
        if (int r = S.get((void**)&t1, br.size() * (size_t)kLvl1Words * 4)) return r;
        B.in0 = (const uint16_t*)in0; B.in1 = (const uint16_t*)in1; B.in2 = (const uint16_t*)in2;
        B.out = t1; B.items = d_br; B.count = (int)br.size();
        if (int r = launch_br(c, B, st)) return r;
        // Not/Copy read their inputs before any key-switch result lands in `out`
        if (int r = launch_affine<uint16_t>((uint16_t*)out, (const uint16_t*)in0, nullptr, d_af, (int)af.size(), kLvl0Words, 0, 0, 0u, st)) return r;
        K.A = t1; K.B = t1; K.out = (uint16_t*)out; K.items = d_ks; K.count = (int)ks.size();
        return launch_ks(c, K, st);
    }
    // level 1: key switch (inputs) -> blind rotate -> final affine into `out`.
    // A Mux needs the second operand from in1 for branch 1 and in2 for branch 0: two
    // key-switch launches, one per source array.
    std::vector<KsItem> ks1, ks2;
    for (const KsItem& k : ks) (k.pad[0] ? ks2 : ks1).push_back(k);
    std::vector<AffItem> afin, afout;
    for (const AffItem& a : af) (a.pad[0] ? afin : afout).push_back(a);
    KsItem *d_k1, *d_k2; AffItem *d_ai, *d_ao; uint16_t* t0; uint32_t* e;
    if (int r = upload(S, &d_k1, ks1)) return r;
    if (int r = upload(S, &d_k2, ks2)) return r;
    if (int r = upload(S, &d_ai, afin)) return r;
    if (int r = upload(S, &d_ao, afout)) return r;
    if (int r = S.get((void**)&t0, ks.size() * (size_t)kLvl0Words * 2)) return r;
    if (int r = S.get((void**)&e, ks.size() * (size_t)kLvl1Words * 4)) return r;
    K.A = (const uint32_t*)in0; K.out = t0;
    K.B = (const uint32_t*)in1; K.items = d_k1; K.count = (int)ks1.size();
    if (int r = launch_ks(c, K, st)) return r;
    K.B = (const uint32_t*)in2; K.items = d_k2; K.count = (int)ks2.size();
    if (int r = launch_ks(c, K, st)) return r;
    B.in0 = t0; B.out = e; B.ca = 1; B.cb = 0; B.off = 0; B.count = (int)ks.size();
    if (int r = launch_br(c, B, st)) return r;
    if (int r = launch_affine<uint32_t>((uint32_t*)out, (const uint32_t*)in0, nullptr, d_ai, (int)afin.size(), kLvl1Words, 0, 0, 0u, st)) return r;
    return launch_affine<uint32_t>((uint32_t*)out, e, e, d_ao, (int)afout.size(), kLvl1Words, 0, 0, 0u, st);
}

}  // namespace

extern "C" {

int b2fhe_gate_batch(int device, b2fhe_stream_t stream, int op, int level, void* out, const void* in0,
                     const void* in1, const void* in2, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(op >= 0 && op < B2FHE_NUM_OPS && (level == 0 || level == 1) && count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out && in0, B2FHE_E_BAD_ARGUMENT);
    return uniform_gate(*c, (cudaStream_t)stream, op, level, out, in0, in1, in2, (int)count);
}

int b2fhe_gate_batch_mixed(int device, b2fhe_stream_t stream, const uint8_t* ops_host, int level, void* out,
                           const void* in0, const void* in1, const void* in2, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE((level == 0 || level == 1) && count < (1u << 29), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(ops_host && out && in0 && in1 && in2, B2FHE_E_BAD_ARGUMENT);
    return mixed_gate(*c, (cudaStream_t)stream, ops_host, level, out, in0, in1, in2, (int)count);
}

int b2fhe_gate_batch_host(int device, b2fhe_stream_t stream, int op, const uint8_t* ops_host, int level,
                          void* out_host, const void* in0_host, const void* in1_host, const void* in2_host,
                          size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE((level == 0 || level == 1) && count < (1u << 29), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_host && in0_host, B2FHE_E_BAD_ARGUMENT);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = count * (level == 0 ? (size_t)kLvl0Words * 2 : (size_t)kLvl1Words * 4);
    Scratch S(st);
    void *d0 = nullptr, *d1 = nullptr, *d2 = nullptr, *dout = nullptr;
    if (int r = S.get(&d0, bytes)) return r;
    if (int r = S.get(&dout, bytes)) return r;
    CK(cudaMemcpyAsync(d0, in0_host, bytes, cudaMemcpyHostToDevice, st));
    if (in1_host) {
        if (int r = S.get(&d1, bytes)) return r;
        CK(cudaMemcpyAsync(d1, in1_host, bytes, cudaMemcpyHostToDevice, st));
    }
    if (in2_host) {
        if (int r = S.get(&d2, bytes)) return r;
        CK(cudaMemcpyAsync(d2, in2_host, bytes, cudaMemcpyHostToDevice, st));
    }
    int r;
    if (ops_host) {
        REQUIRE(d1 && d2, B2FHE_E_BAD_ARGUMENT);
        r = mixed_gate(*c, st, ops_host, level, dout, d0, d1, d2, (int)count);
    } else {
        REQUIRE(op >= 0 && op < B2FHE_NUM_OPS, B2FHE_E_BAD_ARGUMENT);
        r = uniform_gate(*c, st, op, level, dout, d0, d1, d2, (int)count);
    }
    if (r) return r;
    CK(cudaMemcpyAsync(out_host, dout, bytes, cudaMemcpyDeviceToHost, st));
    return 0;
}

int b2fhe_blind_rotate_batch(int device, b2fhe_stream_t stream, uint32_t* out_lvl1, const uint16_t* in_lvl0,
                             uint32_t mu, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_lvl1 && in_lvl0, B2FHE_E_BAD_ARGUMENT);
    BrParams B{};
    B.in0 = in_lvl0; B.out = out_lvl1; B.count = (int)count; B.ca = 1; B.cb = 0; B.off = 0; B.mu = mu;
    return launch_br(*c, B, (cudaStream_t)stream);
}

int b2fhe_bootstrap_tlwe2trlwe_batch(int device, b2fhe_stream_t stream, uint32_t* out_trlwe, const uint16_t* in_lvl0,
                                     uint32_t mu, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_trlwe && in_lvl0, B2FHE_E_BAD_ARGUMENT);
    BrParams B{};
    B.in0 = in_lvl0; B.out = out_trlwe; B.count = (int)count; B.ca = 1; B.cb = 0; B.off = 0; B.mu = mu; B.trlwe = 1;
    return launch_br(*c, B, (cudaStream_t)stream);
}

int b2fhe_bootstrap_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0, const uint16_t* in_lvl0, uint32_t mu,
                          size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_lvl0 && in_lvl0, B2FHE_E_BAD_ARGUMENT);
    cudaStream_t st = (cudaStream_t)stream;
    Scratch S(st);
    uint32_t* t1;
    if (int r = S.get((void**)&t1, count * kLvl1Words * 4)) return r;
    BrParams B{};
    B.in0 = in_lvl0; B.out = t1; B.count = (int)count; B.ca = 1; B.cb = 0; B.off = 0; B.mu = mu;
    if (int r = launch_br(*c, B, st)) return r;
    KsParams K{};
    K.A = t1; K.out = out_lvl0; K.count = (int)count; K.ca = 1; K.cb = 0; K.off = 0u;
    return launch_ks(*c, K, st);
}

int b2fhe_sample_extract_keyswitch_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0, const uint32_t* in_trlwe,
                                         size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_lvl0 && in_trlwe, B2FHE_E_BAD_ARGUMENT);
    KsParams K{};
    K.A = in_trlwe; K.out = out_lvl0; K.count = (int)count; K.ca = 1; K.cb = 0; K.off = 0u; K.trlwe = 1;
    return launch_ks(*c, K, (cudaStream_t)stream);
}

int b2fhe_refresh_batch(int device, b2fhe_stream_t stream, uint32_t* out_trlwe, const uint32_t* in_trlwe, uint32_t mu,
                        size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_trlwe && in_trlwe, B2FHE_E_BAD_ARGUMENT);
    cudaStream_t st = (cudaStream_t)stream;
    Scratch scratch(st);
    void* t0 = nullptr;
    if (int r = scratch.get(&t0, count * kLvl0Words * sizeof(uint16_t))) return r;
    KsParams K{};
    K.A = in_trlwe; K.out = (uint16_t*)t0; K.count = (int)count; K.ca = 1; K.cb = 0; K.off = 0u; K.trlwe = 1;
    if (int r = launch_ks(*c, K, st)) return r;
    BrParams B{};
    B.in0 = (const uint16_t*)t0; B.out = out_trlwe; B.count = (int)count; B.ca = 1; B.cb = 0; B.off = 0; B.mu = mu; B.trlwe = 1;
    return launch_br(*c, B, st);      // ~Scratch frees t0 in stream order
}

int b2fhe_trgsw2ntt_batch(int device, b2fhe_stream_t stream, uint32_t* out_ntt, const uint32_t* in_raw, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 24), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_ntt && in_raw, B2FHE_E_BAD_ARGUMENT);
    const int npoly = (int)count * kRows * 2;
    k_ntt_fwd_s<1><<<ntt_grid(*c, npoly, kNttTeams / 2, 2), kNttThreadsS, ntt_fwd_smem_bytes<kNttTeams>(), (cudaStream_t)stream>>>(
        in_raw, out_ntt, npoly, c->tabs->pr);
    LAUNCHED();
    return 0;
}

int b2fhe_cmux_ntt_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* cs_ntt, const uint32_t* c1,
                         const uint32_t* c0, size_t count, int cs_per_item)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out && cs_ntt && c1 && c0, B2FHE_E_BAD_ARGUMENT);
    CmuxParams P{out, cs_ntt, c1, c0, (int)count, cs_per_item ? 1 : 0};
    cudaStream_t st = (cudaStream_t)stream;
    if ((int)count >= 4 * c->sm_count) k_cmux_s<4><<<((int)count + 3) / 4, 512, cmux_smem_bytes(4), st>>>(P, reinterpret_cast<const PrimeSTables*>(c->tabs));
    else k_cmux_s<1><<<(int)count, 128, cmux_smem_bytes(1), st>>>(P, reinterpret_cast<const PrimeSTables*>(c->tabs));
    LAUNCHED();
    return 0;
}

int b2fhe_keyswitch_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0, const uint32_t* in_lvl1, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (count == 0) return 0;
    REQUIRE(out_lvl0 && in_lvl1, B2FHE_E_BAD_ARGUMENT);
    KsParams K{};
    K.A = in_lvl1; K.out = out_lvl0; K.count = (int)count; K.ca = 1; K.cb = 0; K.off = 0u;
    return launch_ks(*c, K, (cudaStream_t)stream);
}

int b2fhe_ntt1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* in, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (count == 0) return 0;
    REQUIRE(out && in && count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    // inputs prefetched by bulk copies (16-byte aligned sources only; any other 4-byte aligned pointer takes the plain loads)
    if ((reinterpret_cast<uintptr_t>(in) & 15) == 0 && g_ntt_prefetch)
        k_ntt_fwd_s<0, true><<<ntt_grid(*c, count, kNttTeams / 2, 2), kNttThreadsS, ntt_fwd_pf_smem_bytes<kNttTeams>(), (cudaStream_t)stream>>>(
            in, out, (int)count, c->tabs->pr);
    else
        k_ntt_fwd_s<0><<<ntt_grid(*c, count, kNttTeams / 2, 2), kNttThreadsS, ntt_fwd_smem_bytes<kNttTeams>(), (cudaStream_t)stream>>>(
            in, out, (int)count, c->tabs->pr);
    LAUNCHED();
    return 0;
}
int b2fhe_intt1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* in, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (count == 0) return 0;
    REQUIRE(out && in && count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    if (g_ntt_prefetch)      // `in` is read as 128-bit vectors either way
        k_ntt_inv_s<true><<<ntt_grid(*c, count, kNttTeams / 2, 2), kNttThreadsS, ntt_inv_pf_smem_bytes<kNttTeams>(), (cudaStream_t)stream>>>(
            in, out, (int)count, c->tabs->pr);
    else
        k_ntt_inv_s<false><<<ntt_grid(*c, count, kNttTeams / 2, 2), kNttThreadsS, ntt_inv_smem_bytes<kNttTeams>(), (cudaStream_t)stream>>>(
            in, out, (int)count, c->tabs->pr);
    LAUNCHED();
    return 0;
}
int b2fhe_polymul1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* a, const uint32_t* b, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (count == 0) return 0;
    REQUIRE(out && a && b && count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    const bool aligned = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
    if (g_ntt_prefetch && aligned)      // five products per CTA (20 warps per SM), operands prefetched by bulk copies
        k_polymul_s<2, 5, true><<<ntt_grid(*c, count, 5, 1), 640, polymul_pf_smem_bytes<2, 5>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    else if (g_ntt_prefetch)            // 4-byte aligned operands: plain loads
        k_polymul_s<2, 5><<<ntt_grid(*c, count, 5, 1), 640, ntt_s_smem_bytes<2, 10>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    else
        k_polymul_s<2><<<ntt_grid(*c, count, 2, 2), 256, ntt_s_smem_bytes<2, 4>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    LAUNCHED();
    return 0;
}
int b2fhe_polymul1024_exact_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* a, const uint32_t* b, size_t count)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (count == 0) return 0;
    REQUIRE(out && a && b && count < (1u << 30), B2FHE_E_BAD_ARGUMENT);
    const bool aligned = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
    if (g_ntt_prefetch && aligned)      // three products per CTA (18 warps per SM), operands prefetched by bulk copies
        k_polymul_s<3, 3, true><<<ntt_grid(*c, count, 3, 1), 576, polymul_pf_smem_bytes<3, 3>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    else if (g_ntt_prefetch)
        k_polymul_s<3, 3><<<ntt_grid(*c, count, 3, 1), 576, ntt_s_smem_bytes<3, 9>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    else
        k_polymul_s<3><<<ntt_grid(*c, count, 2, 1), 384, ntt_s_smem_bytes<3, 6>(), (cudaStream_t)stream>>>(a, b, out, (int)count, c->tabs->pr);
    LAUNCHED();
    return 0;
}

int b2fhe_int_peak(int device, int mode, double* giga_ops_per_s)
{
    // measured integer-pipe ceiling: mode 0 -> 1e9 IMAD lane-ops/s, mode 1 -> 1e9 butterflies/s
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(giga_ops_per_s && (mode == 0 || mode == 1), B2FHE_E_BAD_ARGUMENT);
    uint32_t* d;
    CK(cudaMalloc(&d, 16));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = 4096, grid = c->sm_count * 8;
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        CK(cudaEventRecord(e0));
        k_int_peak<<<grid, 256>>>(d, iters, mode, 12345u + rep);
        LAUNCHED();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double ops = (double)grid * 256 * iters * (mode == 0 ? 16 : 8);
        const double g = ops / (ms * 1e-3) * 1e-9;
        if (rep > 0 && g > best) best = g;
    }
    *giga_ops_per_s = best;
    CK(cudaEventDestroy(e0));
    CK(cudaEventDestroy(e1));
    CK(cudaFree(d));
    return 0;
}

}  // extern "C"
