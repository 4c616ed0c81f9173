// minnow_ext.inl -- host side of the callers either side of the hot path
// (included by minnow_cuda.cu): the context pool behind CAlloc / DAlloc, the six
// plugin functions, the quantiser alone (quant_QField / quant_Field), integer
// fields, checksums and the segment image.

// ---------------------------------------------------------------------------
// context pool: the reference makes one Compressor / Decompressor per field and
// call (src/funcs.c:122-189); a context costs device allocations, streams, events
// and a jump table, so CFree / DFree park it and CAlloc / DAlloc take it back.
// ---------------------------------------------------------------------------
namespace {
std::mutex g_poolMu;
std::vector<MnwCudaCtx *> g_pool;
constexpr size_t kPoolMax = 8;

// dither seed of the payload a QField was made from (MnwCudaDecompress_v0_9), by
// Quant pointer, until mnwcuda_DequantizeField or mnwcuda_ForgetQuant takes it
std::mutex g_seedMu;
std::unordered_map<const void *, uint64_t> g_quantSeed;

inline uint32_t bits_of(float v) { uint32_t b; memcpy(&b, &v, 4); return b; }
inline float float_of(uint32_t b) { float v; memcpy(&v, &b, 4); return v; }
inline uint32_t bit_length64(uint64_t v) { uint32_t n = 0; while (v) { n++; v >>= 1; } return n; }
inline int32_t run_of(uint32_t fieldCode, int32_t particleLen) {  // values per component of the payload
    return fieldCode == field_Unsi ? (int32_t)(((int64_t)particleLen + 2) / 3) : particleLen;
}
inline unsigned grid_for(int64_t n, int threads) {
    int64_t g = (n + threads - 1) / threads;
    return (unsigned)(g < 1 ? 1 : g > 4096 ? 4096 : g);
}
}  // namespace

extern "C" MnwCudaCtx *mnwcuda_Acquire(int device) {
    if (device < 0) {
        int count = 0;
        if (cudaGetDeviceCount(&count) == cudaSuccess && count > 0) CUDA_CHECK(cudaGetDevice(&device));
    }
    {
        std::lock_guard<std::mutex> lock(g_poolMu);
        for (size_t i = 0; i < g_pool.size(); i++)
            if (g_pool[i]->device == device) {
                MnwCudaCtx *ctx = g_pool[i];
                g_pool.erase(g_pool.begin() + (long)i);
                return ctx;
            }
    }
    return mnwcuda_Create(device);
}

extern "C" void mnwcuda_Release(MnwCudaCtx *ctx) {
    if (!ctx) return;
    ctx->ditherSeed = 0;
    ctx->timing = false;
    {
        std::lock_guard<std::mutex> lock(g_poolMu);
        if (g_pool.size() < kPoolMax) {
            g_pool.push_back(ctx);
            return;
        }
    }
    mnwcuda_Destroy(ctx);
}

extern "C" void *MnwCudaCAlloc_v0_9(void) { return mnwcuda_Acquire(-1); }
extern "C" void MnwCudaCFree_v0_9(void *buffer) { mnwcuda_Release(static_cast<MnwCudaCtx *>(buffer)); }
extern "C" void *MnwCudaDAlloc_v0_9(void) { return mnwcuda_Acquire(-1); }
extern "C" void MnwCudaDFree_v0_9(void *buffer) { mnwcuda_Release(static_cast<MnwCudaCtx *>(buffer)); }

extern "C" void mnwcuda_ForgetQuant(const void *quant) {
    std::lock_guard<std::mutex> lock(g_seedMu);
    g_quantSeed.erase(quant);
}

// ---------------------------------------------------------------------------
// quantised values on the device -> payload (the CFunc body), and back
// ---------------------------------------------------------------------------
namespace {

// d: ParticleLen = values per component (the run length for Unsi), Width = header Width slot.
// d_q: 3 * d.ParticleLen uint64 on the device (not in ioB / ioC). dq: X0, X1, depth for the header.
CField compress_dev(MnwCudaCtx *ctx, FieldHeader hd, const MnwFieldDesc &d, const DevQuant &dq, const uint64_t *d_q,
                    const char *who) {
    cudaStream_t s = ctx->own;
    Plan pl;
    if (build_plan(&d, 1, pl)) MNW_PANIC("%s: bad field.", who);
    const int64_t cap = mnwcuda_FieldBound(d.ParticleLen);
    uint8_t *d_out = static_cast<uint8_t *>(ctx->ioB.need((size_t)cap));
    uint64_t *d_off = static_cast<uint64_t *>(ctx->ioC.need(2 * sizeof(uint64_t)));
    EncodeParams p = prepare_encode(ctx, pl, d_out, cap, d_off, s);
    CUDA_CHECK(cudaMemcpyAsync(p.quant, &dq, sizeof(dq), cudaMemcpyHostToDevice, s));
    if (pl.nBlocks > 0) {
        p.nGroups = (pl.nBlocks + kGroupSlots - 1) / kGroupSlots;
        p.qIn = d_q;
        launch_encode(ctx, p, pl.floatEnd, true, s);
    }
    launch_finalize(ctx, pl, p, s);
    panic_on_status(ctx, s, who);
    CField cf;
    memset(&cf, 0, sizeof(cf));
    cf.Hd = hd;
    cf.Data = fetch_payload(ctx, d_out, d_off, &cf.DataLen);
    return cf;
}

// payload (host) -> 3 * run quantised uint64 in ctx->ioA; h = the payload's header
uint64_t *decompress_dev(MnwCudaCtx *ctx, const CField &cf, MnwDiffHeader &h, int32_t &run, const char *who) {
    run = run_of(cf.Hd.FieldCode, cf.Hd.ParticleLen);
    if (cf.Hd.ParticleLen < 0 || cf.DataLen < mnwcuda_FieldFixedBytes(run)) MNW_PANIC("%s: payload too short.", who);
    cudaStream_t s = ctx->own;
    memcpy(&h, cf.Data, sizeof(h));
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    d.ParticleLen = run;
    d.FieldCode = cf.Hd.FieldCode;
    Plan pl;
    if (build_plan(&d, 1, pl)) MNW_PANIC("%s: bad field.", who);
    const size_t outBytes = (size_t)run * 3 * sizeof(uint64_t);
    uint8_t *d_in = static_cast<uint8_t *>(ctx->ioB.need((size_t)cf.DataLen + 16));
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need(outBytes + 16));
    uint64_t *d_off = static_cast<uint64_t *>(ctx->ioC.need(2 * sizeof(uint64_t)));
    uint64_t off[2] = {0, (uint64_t)cf.DataLen};
    CUDA_CHECK(cudaMemcpyAsync(d_in, cf.Data, (size_t)cf.DataLen, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_off, off, sizeof(off), cudaMemcpyHostToDevice, s));
    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.in = d_in;
    p.offsets = d_off;
    p.fields = static_cast<DevField *>(ctx->fields.need(sizeof(DevField)));
    p.nFields = 1;
    p.uniformTasks = pl.uniformTasks;
    p.quant = static_cast<DevQuant *>(ctx->quant.need(sizeof(DevQuant)));
    p.seeds = static_cast<uint64_t *>(ctx->seeds.need(sizeof(uint64_t)));
    p.tableOf = static_cast<int *>(ctx->tableOf.need(sizeof(int)));
    p.maxStreams = pl.maxStreams;
    p.jumpPow = ctx->jumpPow;
    p.qOut = d_q;
    p.error = ctx->error;
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(p.fields), pl.fields.data(), sizeof(DevField),
                               cudaMemcpyHostToDevice, s));
    k_parse_headers<<<1, 128, 0, s>>>(p);
    ctx->launches++;
    if (pl.nTasks > 0) {
        p.dmgCount = reinterpret_cast<uint32_t *>(ctx->error) + 32;
        p.dmgList = static_cast<uint32_t *>(ctx->dmg.need((size_t)pl.nTasks * 4));
        k_unpack_dequant<true><<<(unsigned)((pl.nTasks + kDecWarps - 1) / kDecWarps), kDecWarps * 32, 0, s>>>(
            p, pl.nTasks);
        k_damaged_tasks<true><<<1, 128, 0, s>>>(p);
        ctx->launches += 2;
    }
    CUDA_CHECK(cudaGetLastError());
    panic_on_status(ctx, s, who);
    return d_q;
}

}  // namespace

// ---------------------------------------------------------------------------
// the plugin functions (src/register.h:15-22): literal QField <-> CField
// ---------------------------------------------------------------------------
namespace {

// header image of a Quantization (what k_finalize writes): DevQuant + the Width slot
void quant_to_header(uint32_t fieldCode, const void *quant, DevQuant &dq, float &widthSlot, const char *who) {
    memset(&dq, 0, sizeof(dq));
    widthSlot = 0.0f;
    if (!quant) MNW_PANIC("%s: QField.Quant is NULL: the quantizer must attach its Quantization.", who);
    if (fieldCode == field_Posn) {
        const PositionQuantization *q = static_cast<const PositionQuantization *>(quant);
        if (q->Depths) MNW_PANIC("Per-particle depths are not supported by algo_Diff v0.9.");
        widthSlot = q->Width;
        for (int c = 0; c < 3; c++) { dq.X0[c] = q->X0[c]; dq.X1[c] = q->X1[c]; }
        dq.depth = q->Depth;
    } else if (fieldCode == field_Velc) {
        const VelocityQuantization *q = static_cast<const VelocityQuantization *>(quant);
        if (q->Depths) MNW_PANIC("Per-particle depths are not supported by algo_Diff v0.9.");
        if (q->SymLog10Scaled) MNW_PANIC("symLog10 not supported.");
        for (int c = 0; c < 3; c++) { dq.X0[c] = q->X0[c]; dq.X1[c] = q->X1[c]; }
        dq.depth = q->Depth;
    } else if (fieldCode == field_Ptid) {
        const IDQuantization *q = static_cast<const IDQuantization *>(quant);
        if (q->Width == 0 || q->Width >= (1ull << 31)) MNW_PANIC("%s: IDQuantization.Width %llu is not supported.", who, (unsigned long long)q->Width);
        uint64_t range = 0;
        for (int c = 0; c < 3; c++) {
            if (q->X1[c] >= (1ull << 32) || q->X1[c] < q->X0[c]) MNW_PANIC("%s: bad IDQuantization.", who);
            dq.X0[c] = float_of((uint32_t)q->X0[c]);
            dq.X1[c] = float_of((uint32_t)q->X1[c]);
            range = std::max(range, q->X1[c] - q->X0[c]);
        }
        if (range >= (1ull << 24)) MNW_PANIC("%s: an id range of %llu grid cells needs more than 24 bits.", who, (unsigned long long)range);
        dq.depth = bit_length64(range);
        widthSlot = float_of((uint32_t)q->Width);
    } else if (fieldCode == field_Unsi) {
        const IntQuantization *q = static_cast<const IntQuantization *>(quant);
        if (q->X1 < q->X0 || q->X1 - q->X0 >= (1ull << 24))
            MNW_PANIC("%s: an integer range of %llu needs more than 24 bits: not supported by algo_Diff v0.9.", who,
                      (unsigned long long)(q->X1 - q->X0));
        dq.X0[0] = float_of((uint32_t)q->X0); dq.X0[1] = float_of((uint32_t)(q->X0 >> 32));
        dq.X1[0] = float_of((uint32_t)q->X1); dq.X1[1] = float_of((uint32_t)(q->X1 >> 32));
        dq.depth = bit_length64(q->X1 - q->X0);
    } else {
        MNW_PANIC("Field code %x is not supported by algo_Diff v0.9.", fieldCode);
    }
    if (dq.depth > 24) MNW_PANIC("Quantization depth %u exceeds 24.", dq.depth);
    if (fieldCode == field_Posn || fieldCode == field_Velc)
        for (int c = 0; c < 3; c++)
            if (dq.maxDiff < dq.X1[c] - dq.X0[c]) dq.maxDiff = dq.X1[c] - dq.X0[c];
}

// malloc'd Quantization from a payload header
void *header_to_quant(uint32_t fieldCode, const MnwDiffHeader &h) {
    if (fieldCode == field_Posn) {
        PositionQuantization *q = static_cast<PositionQuantization *>(calloc(1, sizeof(*q)));
        q->Width = h.Width;
        for (int c = 0; c < 3; c++) { q->X0[c] = h.X0[c]; q->X1[c] = h.X1[c]; }
        q->Depth = h.Depth;
        return q;
    }
    if (fieldCode == field_Velc) {
        VelocityQuantization *q = static_cast<VelocityQuantization *>(calloc(1, sizeof(*q)));
        for (int c = 0; c < 3; c++) { q->X0[c] = h.X0[c]; q->X1[c] = h.X1[c]; }
        q->Depth = h.Depth;
        return q;
    }
    if (fieldCode == field_Ptid) {
        IDQuantization *q = static_cast<IDQuantization *>(calloc(1, sizeof(*q)));
        q->Width = bits_of(h.Width);
        for (int c = 0; c < 3; c++) { q->X0[c] = bits_of(h.X0[c]); q->X1[c] = bits_of(h.X1[c]); }
        return q;
    }
    IntQuantization *q = static_cast<IntQuantization *>(calloc(1, sizeof(*q)));
    q->X0 = (uint64_t)bits_of(h.X0[0]) | ((uint64_t)bits_of(h.X0[1]) << 32);
    q->X1 = (uint64_t)bits_of(h.X1[0]) | ((uint64_t)bits_of(h.X1[1]) << 32);
    return q;
}

}  // namespace

extern "C" CField MnwCudaCompress_v0_9(QField qf, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("MnwCudaCompress_v0_9 needs the buffer returned by MnwCudaCAlloc_v0_9.");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->own;
    const uint32_t code = qf.Hd.FieldCode;
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    DevQuant dq;
    quant_to_header(code, qf.Quant, dq, d.Width, "MnwCudaCompress_v0_9");
    const int32_t n = qf.Hd.ParticleLen;
    if (n < 0) MNW_PANIC("MnwCudaCompress_v0_9: negative ParticleLen.");
    const int32_t run = run_of(code, n);
    d.ParticleLen = run;
    d.FieldCode = code;
    d.DitherSeed = (code == field_Posn || code == field_Velc) ? ctx->ditherSeed : 0;
    // QField.Data: 3N values (Posn, Velc, Ptid) or N values (Unsi, padded here to three runs)
    const size_t given = (size_t)(code == field_Unsi ? n : 3 * (int64_t)n) * sizeof(uint64_t);
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need((size_t)run * 3 * sizeof(uint64_t) + 16));
    if (given) CUDA_CHECK(cudaMemcpyAsync(d_q, qf.Data, given, cudaMemcpyHostToDevice, s));
    if (code == field_Unsi && n > 0 && 3 * (int64_t)run > n) {
        k_u_pad<<<1, 256, 0, s>>>(d_q, n, run);
        ctx->launches++;
    }
    return compress_dev(ctx, qf.Hd, d, dq, d_q, "MnwCudaCompress_v0_9");
}

extern "C" QField MnwCudaDecompress_v0_9(CField cf, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("MnwCudaDecompress_v0_9 needs the buffer returned by MnwCudaDAlloc_v0_9.");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    MnwDiffHeader h;
    int32_t run = 0;
    const uint64_t *d_q = decompress_dev(ctx, cf, h, run, "MnwCudaDecompress_v0_9");
    const uint32_t code = cf.Hd.FieldCode;
    const size_t outBytes = (size_t)(code == field_Unsi ? cf.Hd.ParticleLen : 3 * (int64_t)cf.Hd.ParticleLen) * sizeof(uint64_t);
    QField qf;
    memset(&qf, 0, sizeof(qf));
    qf.Hd = cf.Hd;
    qf.Data = static_cast<uint64_t *>(xmalloc(outBytes));
    if (outBytes) CUDA_CHECK(cudaMemcpyAsync(qf.Data, d_q, outBytes, cudaMemcpyDeviceToHost, ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    qf.Quant = header_to_quant(code, h);
    if (code == field_Posn || code == field_Velc) {
        std::lock_guard<std::mutex> lock(g_seedMu);
        g_quantSeed[qf.Quant] = h.DitherSeed;
    }
    return qf;
}

// ---------------------------------------------------------------------------
// the quantiser alone
// ---------------------------------------------------------------------------
namespace {

// Posn / Velc floats (host) -> 3n quantised uint64 in ctx->ioA; dq = the parameters found
uint64_t *quantize_floats_dev(MnwCudaCtx *ctx, const Field &f, DevQuant &dq, float &delta, float &width) {
    cudaStream_t s = ctx->own;
    accuracy_of(f, delta, width);
    const int32_t n = f.Hd.ParticleLen;
    if (n < 0) MNW_PANIC("quantizer: negative ParticleLen.");
    const size_t inBytes = (size_t)n * 3 * sizeof(float);
    float *d_in = static_cast<float *>(ctx->ioD.need(inBytes + 16));
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need((size_t)n * 3 * sizeof(uint64_t) + 16));
    if (inBytes) CUDA_CHECK(cudaMemcpyAsync(d_in, f.Data, inBytes, cudaMemcpyHostToDevice, s));
    DevField fd;
    memset(&fd, 0, sizeof(fd));
    fd.n = n;
    fd.nb = (int32_t)nblocks_of(n);
    fd.fieldCode = f.Hd.FieldCode;
    fd.delta = delta;
    fd.width = width;
    EncodeParams ep;
    memset(&ep, 0, sizeof(ep));
    ep.fields = static_cast<DevField *>(ctx->fields.need(sizeof(DevField)));
    ep.nFields = 1;
    ep.minmax = static_cast<uint32_t *>(ctx->minmax.need(6 * sizeof(uint32_t)));
    ep.quant = static_cast<DevQuant *>(ctx->quant.need(sizeof(DevQuant)));
    ep.error = ctx->error;
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(ep.fields), &fd, sizeof(fd), cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemsetAsync(ep.minmax, 0xff, 6 * sizeof(uint32_t), s));
    QuantParams qp;
    qp.in = d_in;
    qp.qOut = d_q;
    qp.n = n;
    qp.isPos = f.Hd.FieldCode == field_Posn;
    qp.L = width;
    qp.minmax = ep.minmax;
    qp.quant = ep.quant;
    const dim3 grid((unsigned)((n + kQThreads * kQPerThread - 1) / (kQThreads * kQPerThread)), 3);
    if (n > 0) k_q_minmax<<<grid, kQThreads, 0, s>>>(qp);
    k_q_params<<<1, 1, 0, s>>>(ep);
    if (n > 0) k_q_quantise<<<grid, kQThreads, 0, s>>>(qp);
    ctx->launches += n > 0 ? 3 : 1;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(&dq, ep.quant, sizeof(dq), cudaMemcpyDeviceToHost, s));
    panic_on_status(ctx, s, "quantizer");  // waits; depth > 24 is the Panic of src/quant.c:688-693
    return d_q;
}

// Ptid ids (host) -> 3n uint64 in ctx->ioA + IDQuantization
uint64_t *quantize_ids_dev(MnwCudaCtx *ctx, const Field &f, IDQuantization &quant) {
    cudaStream_t s = ctx->own;
    const IDAccuracy *acc = static_cast<const IDAccuracy *>(f.Acc);
    if (!acc) MNW_PANIC("Field has no IDAccuracy.");
    if (acc->Width == 0 || acc->Width >= (1ull << 31)) MNW_PANIC("IDAccuracy.Width %llu is not supported.", (unsigned long long)acc->Width);
    const int32_t n = f.Hd.ParticleLen;
    if (n < 0) MNW_PANIC("quantizer: negative ParticleLen.");
    memset(&quant, 0, sizeof(quant));
    quant.Width = acc->Width;
    uint64_t *d_ids = static_cast<uint64_t *>(ctx->ioD.need((size_t)n * 8 + 64));
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need((size_t)n * 24 + 16));
    if (n == 0) return d_q;
    long long *d_mm = reinterpret_cast<long long *>(ctx->minmax.need(6 * sizeof(long long)));
    long long init[6] = {LLONG_MAX, LLONG_MIN, LLONG_MAX, LLONG_MIN, LLONG_MAX, LLONG_MIN}, mm[6];
    CUDA_CHECK(cudaMemcpyAsync(d_ids, f.Data, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_mm, init, sizeof(init), cudaMemcpyHostToDevice, s));
    IdParams ip;
    ip.ids = d_ids;
    ip.q = d_q;
    ip.n = n;
    ip.width = acc->Width;
    ip.mm = d_mm;
    k_id_minmax<<<grid_for(n, 256), 256, 0, s>>>(ip);
    k_id_quantise<<<grid_for(n, 256), 256, 0, s>>>(ip);
    ctx->launches += 2;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    for (int c = 0; c < 3; c++) {  // util_U64UndoPeriodic shifts a component with a negative minimum up by L
        const long long shift = mm[2 * c] < 0 ? (long long)acc->Width : 0;
        quant.X0[c] = (uint64_t)(mm[2 * c] + shift);
        quant.X1[c] = (uint64_t)(mm[2 * c + 1] + shift);
    }
    return d_q;
}

// Unsi values (host) -> 3*run uint64 in ctx->ioA (tail padded) + IntQuantization
uint64_t *quantize_uints_dev(MnwCudaCtx *ctx, const Field &f, IntQuantization &quant) {
    cudaStream_t s = ctx->own;
    const int32_t n = f.Hd.ParticleLen;
    if (n < 0) MNW_PANIC("quantizer: negative ParticleLen.");
    const int32_t run = run_of(field_Unsi, n);
    memset(&quant, 0, sizeof(quant));
    uint64_t *d_data = static_cast<uint64_t *>(ctx->ioD.need((size_t)n * 8 + 64));
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need((size_t)run * 24 + 16));
    if (n == 0) return d_q;
    unsigned long long *d_mm = reinterpret_cast<unsigned long long *>(ctx->minmax.need(2 * sizeof(unsigned long long)));
    unsigned long long init[2] = {~0ull, 0ull}, mm[2];
    CUDA_CHECK(cudaMemcpyAsync(d_data, f.Data, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_mm, init, sizeof(init), cudaMemcpyHostToDevice, s));
    k_u_minmax<<<grid_for(n, 256), 256, 0, s>>>(d_data, n, d_mm);
    k_u_quantise<<<grid_for(3 * (int64_t)run, 256), 256, 0, s>>>(d_data, n, run, d_mm, d_q);
    ctx->launches += 2;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    quant.X0 = mm[0];
    quant.X1 = mm[1];
    return d_q;
}

}  // namespace

extern "C" QField mnwcuda_QuantizeField(Field f, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("mnwcuda_QuantizeField needs a context (MnwCudaCAlloc_v0_9 / mnwcuda_Acquire).");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    const uint32_t code = f.Hd.FieldCode;
    const int32_t n = f.Hd.ParticleLen;
    QField qf;
    memset(&qf, 0, sizeof(qf));
    qf.Hd = f.Hd;
    qf.Valid = 1;
    const uint64_t *d_q = nullptr;
    size_t count = (size_t)(n > 0 ? n : 0) * 3;
    if (code == field_Posn || code == field_Velc) {
        DevQuant dq;
        float delta, width;
        d_q = quantize_floats_dev(ctx, f, dq, delta, width);
        if (code == field_Posn) {
            const PositionAccuracy *acc = static_cast<const PositionAccuracy *>(f.Acc);
            PositionQuantization *q = static_cast<PositionQuantization *>(calloc(1, sizeof(*q)));
            q->Len = acc->Len;
            q->Width = acc->Width;
            for (int c = 0; c < 3; c++) { q->X0[c] = dq.X0[c]; q->X1[c] = dq.X1[c]; }
            q->Depth = (uint8_t)dq.depth;
            qf.Quant = q;
        } else {
            const VelocityAccuracy *acc = static_cast<const VelocityAccuracy *>(f.Acc);
            VelocityQuantization *q = static_cast<VelocityQuantization *>(calloc(1, sizeof(*q)));
            q->Len = acc->Len;
            q->SymLog10Threshold = acc->SymLog10Threshold;
            for (int c = 0; c < 3; c++) { q->X0[c] = dq.X0[c]; q->X1[c] = dq.X1[c]; }
            q->Depth = (uint8_t)dq.depth;
            qf.Quant = q;
        }
    } else if (code == field_Ptid) {
        IDQuantization *q = static_cast<IDQuantization *>(calloc(1, sizeof(*q)));
        d_q = quantize_ids_dev(ctx, f, *q);
        qf.Quant = q;
    } else if (code == field_Unsi) {
        IntQuantization *q = static_cast<IntQuantization *>(calloc(1, sizeof(*q)));
        d_q = quantize_uints_dev(ctx, f, *q);
        qf.Quant = q;
        count = (size_t)(n > 0 ? n : 0);  // uint() returns N values (src/quant.c:381)
    } else {
        MNW_PANIC("Unrecognized field code %x.", code);  // src/quant.c:142
    }
    qf.Data = static_cast<uint64_t *>(xmalloc(count * sizeof(uint64_t)));
    if (count) CUDA_CHECK(cudaMemcpyAsync(qf.Data, d_q, count * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    return qf;
}

namespace {

// 3n quantised uint64 on the device -> dithered floats (host, malloc'd): quant_Field for Posn / Velc
float *dequantize_floats(MnwCudaCtx *ctx, uint32_t code, int32_t n, const DevQuant &dqIn, float widthSlot,
                         uint64_t seed, const uint64_t *d_q) {
    cudaStream_t s = ctx->own;
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    d.ParticleLen = n;
    d.FieldCode = code;
    Plan pl;
    if (build_plan(&d, 1, pl)) MNW_PANIC("dequantizer: bad field.");
    const size_t outBytes = (size_t)n * 3 * sizeof(float);
    float *out = static_cast<float *>(xmalloc(outBytes));
    if (n == 0) return out;
    DevQuant dq = dqIn;
    dq.width = widthSlot;
    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.fields = static_cast<DevField *>(ctx->fields.need(sizeof(DevField)));
    p.nFields = 1;
    p.quant = static_cast<DevQuant *>(ctx->quant.need(sizeof(DevQuant)));
    p.seeds = static_cast<uint64_t *>(ctx->seeds.need(sizeof(uint64_t)));
    p.tableOf = static_cast<int *>(ctx->tableOf.need(sizeof(int)));
    p.maxStreams = pl.maxStreams;
    p.states = static_cast<uint64_t *>(ctx->states.need((size_t)pl.maxStreams * 16 + 16));
    p.jumpPow = ctx->jumpPow;
    p.out = static_cast<float *>(ctx->ioD.need(outBytes + 16));
    p.error = ctx->error;
    const int zero = 0;
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(p.fields), pl.fields.data(), sizeof(DevField),
                               cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(p.quant, &dq, sizeof(dq), cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(p.seeds, &seed, sizeof(seed), cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(p.tableOf, &zero, sizeof(zero), cudaMemcpyHostToDevice, s));
    k_stream_states<<<dim3((unsigned)((pl.maxStreams + 127) / 128), 1), 128, 0, s>>>(p);
    k_dequant_q<<<(unsigned)((pl.nTasks + 3) / 4), 128, 0, s>>>(p, d_q, pl.nTasks);
    ctx->launches += 2;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(out, p.out, outBytes, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    return out;
}

uint64_t *restore_ids(MnwCudaCtx *ctx, int32_t n, const IDQuantization &q, const uint64_t *d_q) {
    cudaStream_t s = ctx->own;
    uint64_t *out = static_cast<uint64_t *>(xmalloc((size_t)n * 8));
    if (n == 0) return out;
    uint64_t *d_ids = static_cast<uint64_t *>(ctx->ioD.need((size_t)n * 8 + 16));
    k_id_restore<<<grid_for(n, 256), 256, 0, s>>>(d_q, n, q.Width, q.X0[0], q.X0[1], q.X0[2], d_ids);
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(out, d_ids, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    return out;
}

uint64_t *restore_uints(MnwCudaCtx *ctx, int32_t n, const IntQuantization &q, const uint64_t *d_q) {
    cudaStream_t s = ctx->own;
    uint64_t *out = static_cast<uint64_t *>(xmalloc((size_t)n * 8));
    if (n == 0) return out;
    uint64_t *d_out = static_cast<uint64_t *>(ctx->ioD.need((size_t)n * 8 + 16));
    k_u_restore<<<grid_for(n, 256), 256, 0, s>>>(d_q, n, q.X0, d_out);
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    return out;
}

// reported accuracy of a float field: depthToDelta on component 0 (src/quant.c:490-493, 654-662)
void *accuracy_from(uint32_t code, const float X0[3], const float X1[3], uint32_t depth, float width) {
    const float delta = (X1[0] - X0[0]) / (float)(1 << depth);
    if (code == field_Posn) {
        PositionAccuracy *a = static_cast<PositionAccuracy *>(calloc(1, sizeof(*a)));
        a->Delta = delta;
        a->Width = width;
        return a;
    }
    VelocityAccuracy *a = static_cast<VelocityAccuracy *>(calloc(1, sizeof(*a)));
    a->Delta = delta;
    return a;
}

}  // namespace

extern "C" Field mnwcuda_DequantizeField(QField qf, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("mnwcuda_DequantizeField needs a context (MnwCudaDAlloc_v0_9 / mnwcuda_Acquire).");
    if (!qf.Quant) MNW_PANIC("mnwcuda_DequantizeField: QField.Quant is NULL.");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->own;
    const uint32_t code = qf.Hd.FieldCode;
    const int32_t n = qf.Hd.ParticleLen;
    if (n < 0) MNW_PANIC("mnwcuda_DequantizeField: negative ParticleLen.");
    Field f;
    memset(&f, 0, sizeof(f));
    f.Hd = qf.Hd;
    f.Valid = 1;
    const size_t count = (size_t)n * (code == field_Unsi ? 1 : 3);
    uint64_t *d_q = static_cast<uint64_t *>(ctx->ioA.need(count * 8 + 16));
    if (count) CUDA_CHECK(cudaMemcpyAsync(d_q, qf.Data, count * 8, cudaMemcpyHostToDevice, s));
    if (code == field_Posn || code == field_Velc) {
        DevQuant dq;
        float widthSlot;
        quant_to_header(code, qf.Quant, dq, widthSlot, "mnwcuda_DequantizeField");
        uint64_t seed = ctx->ditherSeed;
        {
            std::lock_guard<std::mutex> lock(g_seedMu);
            auto it = g_quantSeed.find(qf.Quant);
            if (it != g_quantSeed.end()) {
                seed = it->second;
                g_quantSeed.erase(it);
            }
        }
        f.Data = dequantize_floats(ctx, code, n, dq, widthSlot, seed, d_q);
        f.Acc = accuracy_from(code, dq.X0, dq.X1, dq.depth, widthSlot);
    } else if (code == field_Ptid) {
        const IDQuantization *q = static_cast<const IDQuantization *>(qf.Quant);
        f.Data = restore_ids(ctx, n, *q, d_q);
        IDAccuracy *a = static_cast<IDAccuracy *>(calloc(1, sizeof(*a)));
        a->Width = q->Width;
        f.Acc = a;
    } else if (code == field_Unsi) {
        f.Data = restore_uints(ctx, n, *static_cast<const IntQuantization *>(qf.Quant), d_q);
        f.Acc = nullptr;  // "No need to set Acc", src/quant.c:604
    } else {
        MNW_PANIC("Unrecognized field code %x.", code);  // src/quant.c:153
    }
    return f;
}

// mnwcuda_EncodeField for Ptid / Unsi: quantiser and CFunc back to back on the device
static CField encode_int_field(Field f, MnwCudaCtx *ctx) {
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    d.FieldCode = f.Hd.FieldCode;
    d.ParticleLen = run_of(f.Hd.FieldCode, f.Hd.ParticleLen);
    DevQuant dq;
    const uint64_t *d_q;
    if (f.Hd.FieldCode == field_Ptid) {
        IDQuantization q;
        d_q = quantize_ids_dev(ctx, f, q);
        quant_to_header(field_Ptid, &q, dq, d.Width, "mnwcuda_EncodeField");
    } else {
        IntQuantization q;
        d_q = quantize_uints_dev(ctx, f, q);
        quant_to_header(field_Unsi, &q, dq, d.Width, "mnwcuda_EncodeField");
    }
    return compress_dev(ctx, f.Hd, d, dq, d_q, "mnwcuda_EncodeField");
}

static Field decode_int_field(CField cf, MnwCudaCtx *ctx) {
    MnwDiffHeader h;
    int32_t run = 0;
    const uint64_t *d_q = decompress_dev(ctx, cf, h, run, "mnwcuda_DecodeField");
    Field f;
    memset(&f, 0, sizeof(f));
    f.Hd = cf.Hd;
    f.Valid = 1;
    void *quant = header_to_quant(cf.Hd.FieldCode, h);
    if (cf.Hd.FieldCode == field_Ptid) {
        const IDQuantization *q = static_cast<const IDQuantization *>(quant);
        f.Data = restore_ids(ctx, cf.Hd.ParticleLen, *q, d_q);
        IDAccuracy *a = static_cast<IDAccuracy *>(calloc(1, sizeof(*a)));
        a->Width = q->Width;
        f.Acc = a;
    } else {
        f.Data = restore_uints(ctx, cf.Hd.ParticleLen, *static_cast<const IntQuantization *>(quant), d_q);
    }
    free(quant);
    return f;
}

// ---------------------------------------------------------------------------
// checksums and the segment image
// ---------------------------------------------------------------------------
extern "C" int64_t mnwcuda_BlockCount(const MnwFieldDesc *h_fields, int64_t nFields) {
    int64_t total = 0;
    for (int64_t i = 0; i < nFields; i++) total += 1 + 3 * nblocks_of(h_fields[i].ParticleLen);
    return total;
}

namespace {
// DecodeParams over a packed stream on the device, for the block-header kernels
int block_params(MnwCudaCtx *ctx, const uint8_t *d_stream, const uint64_t *d_offsets, const MnwFieldDesc *h_fields,
                 int64_t nFields, DecodeParams &p, int64_t &nEntries, cudaStream_t s) {
    Plan pl;
    int rc = build_plan(h_fields, nFields, pl);
    if (rc) return rc;
    memset(&p, 0, sizeof(p));
    p.in = d_stream;
    p.offsets = d_offsets;
    p.fields = static_cast<DevField *>(ctx->fields.need((size_t)nFields * sizeof(DevField)));
    p.nFields = (int)nFields;
    p.uniformTasks = pl.uniformBlocks;  // k_block_headers reads it as blocks per field (0: fields differ)
    p.error = ctx->error;
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(p.fields), pl.fields.data(), (size_t)nFields * sizeof(DevField),
                               cudaMemcpyHostToDevice, s));
    nEntries = pl.nBlocks + nFields;
    return 0;
}
}  // namespace

extern "C" int mnwcuda_BlockHeadersDev(MnwCudaCtx *ctx, const uint8_t *d_stream, const uint64_t *d_offsets,
                                       const MnwFieldDesc *h_fields, int64_t nFields, MnwBlockHeader *d_headers,
                                       void *stream) {
    if (!ctx || nFields < 0) return -1;
    if (nFields == 0) return 0;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DecodeParams p;
    int64_t nEntries = 0;
    int rc = block_params(ctx, d_stream, d_offsets, h_fields, nFields, p, nEntries, s);
    if (rc) return rc;
    mark(ctx, 3, s);
    k_block_headers<<<(unsigned)(nFields + (nEntries - nFields + 127) / 128), 128, 0, s>>>(
        p, reinterpret_cast<MnwBlockHeaderDev *>(d_headers), 0, nullptr, nEntries);
    mark(ctx, 4, s);
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int mnwcuda_VerifyBlocksDev(MnwCudaCtx *ctx, uint8_t *d_stream, const uint64_t *d_offsets,
                                       const MnwFieldDesc *h_fields, int64_t nFields, const MnwBlockHeader *d_headers,
                                       int32_t *d_fieldValid, void *stream) {
    if (!ctx || nFields < 0) return -1;
    if (nFields == 0) return 0;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DecodeParams p;
    int64_t nEntries = 0;
    int rc = block_params(ctx, d_stream, d_offsets, h_fields, nFields, p, nEntries, s);
    if (rc) return rc;
    uint8_t *d_damaged = static_cast<uint8_t *>(ctx->misc.need((size_t)nEntries + 16));
    std::vector<int32_t> ones((size_t)nFields, 1);
    CUDA_CHECK(cudaMemcpyAsync(d_fieldValid, ones.data(), (size_t)nFields * 4, cudaMemcpyHostToDevice, s));
    const unsigned grid = (unsigned)((nEntries + 127) / 128);
    k_block_headers<<<(unsigned)(nFields + (nEntries - nFields + 127) / 128), 128, 0, s>>>(
        p, reinterpret_cast<MnwBlockHeaderDev *>(const_cast<MnwBlockHeader *>(d_headers)), 1, d_damaged, nEntries);
    k_mark_damage<<<grid, 128, 0, s>>>(p, d_damaged, d_fieldValid, nEntries);
    ctx->launches += 2;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(s));  // `ones` is pageable and must outlive the copy
    return 0;
}

extern "C" int mnwcuda_ChecksumsHost(MnwCudaCtx *ctx, const uint8_t *h_data, int64_t dataLen, const int64_t *h_off,
                                     const int64_t *h_len, int64_t nRanges, uint32_t *h_sums) {
    if (!ctx || dataLen < 0 || nRanges < 0) return -1;
    if (nRanges == 0) return 0;
    for (int64_t r = 0; r < nRanges; r++)
        if (h_off[r] < 0 || h_len[r] < 0 || h_off[r] + h_len[r] > dataLen) return -1;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->own;
    uint8_t *d_data = static_cast<uint8_t *>(ctx->ioB.need((size_t)dataLen + 16));
    int64_t *d_rng = static_cast<int64_t *>(ctx->ioC.need((size_t)nRanges * 20 + 64));
    uint32_t *d_sums = reinterpret_cast<uint32_t *>(d_rng + 2 * nRanges);
    if (dataLen) CUDA_CHECK(cudaMemcpyAsync(d_data, h_data, (size_t)dataLen, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_rng, h_off, (size_t)nRanges * 8, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_rng + nRanges, h_len, (size_t)nRanges * 8, cudaMemcpyHostToDevice, s));
    k_checksum_ranges<<<(unsigned)((nRanges + 127) / 128), 128, 0, s>>>(d_data, d_rng, d_rng + nRanges, d_sums, nRanges);
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(h_sums, d_sums, (size_t)nRanges * 4, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    return 0;
}

namespace {
bool seg_desc(const FieldHeader &hd, MnwFieldDesc &d) {
    memset(&d, 0, sizeof(d));
    if (hd.ParticleLen < 0) return false;
    d.ParticleLen = run_of(hd.FieldCode, hd.ParticleLen);
    d.FieldCode = hd.FieldCode;
    return hd.FieldCode == field_Posn || hd.FieldCode == field_Velc || hd.FieldCode == field_Ptid ||
           hd.FieldCode == field_Unsi;
}
}  // namespace

extern "C" int64_t mnwcuda_SegmentImageBound(const CField *fields, int32_t nFields) {
    int64_t total = 16 + 16 * (int64_t)nFields;
    for (int32_t i = 0; i < nFields; i++) {
        MnwFieldDesc d;
        if (!seg_desc(fields[i].Hd, d)) return -3;
        total += 8 * (1 + 3 * nblocks_of(d.ParticleLen)) + fields[i].DataLen;
    }
    return total;
}

extern "C" int64_t mnwcuda_SegmentToImage(MnwCudaCtx *ctx, const CField *fields, int32_t nFields,
                                          int32_t particleNum, uint8_t *h_image, int64_t cap) {
    if (!ctx || nFields < 0) return -1;
    const int64_t bound = mnwcuda_SegmentImageBound(fields, nFields);
    if (bound < 0) return bound;
    if (bound > cap) return -2;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->own;
    std::vector<MnwFieldDesc> descs((size_t)nFields);
    std::vector<uint64_t> offs((size_t)nFields + 1, 0);
    for (int32_t i = 0; i < nFields; i++) {
        seg_desc(fields[i].Hd, descs[(size_t)i]);
        if (fields[i].DataLen < mnwcuda_FieldFixedBytes(descs[(size_t)i].ParticleLen) || (fields[i].DataLen & 15)) return -1;
        offs[(size_t)i + 1] = offs[(size_t)i] + (uint64_t)fields[i].DataLen;
    }
    const int64_t nBlocks = mnwcuda_BlockCount(descs.data(), nFields);
    if (nBlocks >= (1ll << 31)) return -5;
    const int64_t hdrBytes = 16 + 16 * (int64_t)nFields + 8 * nBlocks;
    // image: headers, then the fields' payloads back to back (= their blocks in order)
    int32_t *sh = reinterpret_cast<int32_t *>(h_image);
    sh[0] = 0;
    sh[1] = (int32_t)nBlocks;
    sh[2] = nFields;
    sh[3] = particleNum;
    for (int32_t i = 0; i < nFields; i++) {
        memcpy(h_image + 16 + 16 * (int64_t)i, &fields[i].Hd, 16);  // src/types.h:103-108, what src/funcs.c:85 writes
        if (fields[i].DataLen) memcpy(h_image + hdrBytes + offs[(size_t)i], fields[i].Data, (size_t)fields[i].DataLen);
    }
    // per-block checksums on the device
    const uint64_t total = offs[(size_t)nFields];
    uint8_t *d_stream = static_cast<uint8_t *>(ctx->ioB.need((size_t)total + 16));
    uint8_t *d_aux = static_cast<uint8_t *>(ctx->ioD.need((size_t)(nFields + 1) * 8 + (size_t)nBlocks * 8 + 64));
    uint64_t *d_off = reinterpret_cast<uint64_t *>(d_aux);
    MnwBlockHeader *d_hdr = reinterpret_cast<MnwBlockHeader *>(d_aux + (size_t)(nFields + 1) * 8);
    if (total) CUDA_CHECK(cudaMemcpyAsync(d_stream, h_image + hdrBytes, (size_t)total, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_off, offs.data(), (size_t)(nFields + 1) * 8, cudaMemcpyHostToDevice, s));
    if (nFields > 0) {
        int rc = mnwcuda_BlockHeadersDev(ctx, d_stream, d_off, descs.data(), nFields, d_hdr, s);
        if (rc) return rc;
        CUDA_CHECK(cudaMemcpyAsync(h_image + 16 + 16 * (int64_t)nFields, d_hdr, (size_t)nBlocks * 8, cudaMemcpyDeviceToHost, s));
    }
    if (mnwcuda_Status(ctx, s) != 0) return -1;  // a table that points outside its payload
    // SegmentHeader.Checksum over the 12 + 16F + 8B bytes behind it (tex:150-158)
    const int64_t o = 4, l = hdrBytes - 4;
    uint32_t top = 0;
    int rc = mnwcuda_ChecksumsHost(ctx, h_image, hdrBytes, &o, &l, 1, &top);
    if (rc) return rc;
    memcpy(h_image, &top, 4);
    return hdrBytes + (int64_t)total;
}

extern "C" int32_t mnwcuda_SegmentFromImage(MnwCudaCtx *ctx, const uint8_t *h_image, int64_t len, CField *fields,
                                            int32_t maxFields, int32_t *valid, int32_t *particleNum) {
    if (!ctx || len < 16) return -1;
    int32_t sh[4];
    memcpy(sh, h_image, 16);
    const int64_t nBlocks = sh[1];
    const int32_t nFields = sh[2];
    if (nFields < 0 || nBlocks < 0 || nFields > maxFields) return -1;
    const int64_t hdrBytes = 16 + 16 * (int64_t)nFields + 8 * nBlocks;
    if (hdrBytes > len) return -1;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->own;
    const int64_t o = 4, l = hdrBytes - 4;
    uint32_t top = 0;
    if (mnwcuda_ChecksumsHost(ctx, h_image, hdrBytes, &o, &l, 1, &top)) return -1;
    if (top != (uint32_t)sh[0]) return -4;
    if (particleNum) *particleNum = sh[3];
    // field extents from the block lengths
    std::vector<MnwFieldDesc> descs((size_t)nFields);
    std::vector<uint64_t> offs((size_t)nFields + 1, 0);
    const MnwBlockHeader *bh = reinterpret_cast<const MnwBlockHeader *>(h_image + 16 + 16 * (int64_t)nFields);
    int64_t seen = 0;
    for (int32_t i = 0; i < nFields; i++) {
        memset(&fields[i], 0, sizeof(fields[i]));
        memcpy(&fields[i].Hd, h_image + 16 + 16 * (int64_t)i, 16);
        if (!seg_desc(fields[i].Hd, descs[(size_t)i])) return -3;
        const int64_t nbk = 1 + 3 * nblocks_of(descs[(size_t)i].ParticleLen);
        if (seen + nbk > nBlocks) return -1;
        int64_t bytes = 0;
        for (int64_t k = 0; k < nbk; k++) {
            if (bh[seen + k].Length < 0 || (bh[seen + k].Length & 15)) return -1;
            bytes += bh[seen + k].Length;
        }
        if (bh[seen].Length != mnwcuda_FieldFixedBytes(descs[(size_t)i].ParticleLen)) return -1;
        offs[(size_t)i + 1] = offs[(size_t)i] + (uint64_t)bytes;
        seen += nbk;
    }
    if (seen != nBlocks || hdrBytes + (int64_t)offs[(size_t)nFields] > len) return -1;
    const uint64_t total = offs[(size_t)nFields];
    uint8_t *d_stream = static_cast<uint8_t *>(ctx->ioB.need((size_t)total + 16));
    uint8_t *d_aux = static_cast<uint8_t *>(
        ctx->ioD.need((size_t)(nFields + 1) * 8 + (size_t)nBlocks * 8 + (size_t)nFields * 4 + 64));
    uint64_t *d_off = reinterpret_cast<uint64_t *>(d_aux);
    MnwBlockHeader *d_hdr = reinterpret_cast<MnwBlockHeader *>(d_aux + (size_t)(nFields + 1) * 8);
    int32_t *d_valid = reinterpret_cast<int32_t *>(d_aux + (size_t)(nFields + 1) * 8 + (size_t)nBlocks * 8);
    if (total) CUDA_CHECK(cudaMemcpyAsync(d_stream, h_image + hdrBytes, (size_t)total, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_off, offs.data(), (size_t)(nFields + 1) * 8, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(d_hdr, bh, (size_t)nBlocks * 8, cudaMemcpyHostToDevice, s));
    if (nFields > 0) {
        int rc = mnwcuda_VerifyBlocksDev(ctx, d_stream, d_off, descs.data(), nFields, d_hdr, d_valid, s);
        if (rc) return rc;
        CUDA_CHECK(cudaMemcpyAsync(valid, d_valid, (size_t)nFields * 4, cudaMemcpyDeviceToHost, s));
    }
    // the payloads come back from the device: damaged blocks are marked in their tables
    std::vector<int64_t> co((size_t)nFields), cl((size_t)nFields);
    for (int32_t i = 0; i < nFields; i++) {
        fields[i].DataLen = (int64_t)(offs[(size_t)i + 1] - offs[(size_t)i]);
        fields[i].Data = static_cast<uint8_t *>(xmalloc((size_t)fields[i].DataLen));
        if (fields[i].DataLen)
            CUDA_CHECK(cudaMemcpyAsync(fields[i].Data, d_stream + offs[(size_t)i], (size_t)fields[i].DataLen,
                                       cudaMemcpyDeviceToHost, s));
        co[(size_t)i] = (int64_t)offs[(size_t)i];
        cl[(size_t)i] = fields[i].DataLen;
    }
    CUDA_CHECK(cudaStreamSynchronize(s));
    // CField.Checksum of what is handed on, so that Decompress (src/funcs.c:45-56) accepts it: one thread per field
    if (nFields > 0) {
        std::vector<uint32_t> sums((size_t)nFields);
        int64_t *d_rng = static_cast<int64_t *>(ctx->ioC.need((size_t)nFields * 20 + 64));
        uint32_t *d_sums = reinterpret_cast<uint32_t *>(d_rng + 2 * (int64_t)nFields);
        CUDA_CHECK(cudaMemcpyAsync(d_rng, co.data(), (size_t)nFields * 8, cudaMemcpyHostToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(d_rng + nFields, cl.data(), (size_t)nFields * 8, cudaMemcpyHostToDevice, s));
        k_checksum_ranges<<<(unsigned)((nFields + 127) / 128), 128, 0, s>>>(d_stream, d_rng, d_rng + nFields, d_sums, nFields);
        ctx->launches++;
        CUDA_CHECK(cudaMemcpyAsync(sums.data(), d_sums, (size_t)nFields * 4, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        // (a field whose header block is damaged must not pass that test)
        for (int32_t i = 0; i < nFields; i++) fields[i].Checksum = valid[i] ? sums[(size_t)i] : ~sums[(size_t)i];
    }
    return nFields;
}
