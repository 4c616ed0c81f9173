// minnow_noalloc.inl -- mnwcuda_DecodeFieldNoAlloc (included by minnow_cuda.cu, so
// the array below and the cudaGetSymbolAddress that resolves it share one
// translation unit). No counterpart in the reference: a decode of one float field
// for callers that may not allocate (CUDA graph capture, serving loops). The
// decoder's scratch -- control words, field record, parsed header, jump-matrix
// powers, damaged-task list, dither stream states -- lives in a file-scope device
// array of fixed size instead of a context's cudaMalloc'd buffers; the kernels are
// the ones mnwcuda_DecodeFieldsDev launches, handed pointers into that array.
// Every GPU that loads the library holds the array until exit.

namespace {

constexpr int32_t kNaMaxParticles = MNWCUDA_NOALLOC_MAX_PARTICLES;
constexpr int64_t kNaMaxBlocks = 3 * (((int64_t)kNaMaxParticles + kBlockLen - 1) / kBlockLen);
constexpr int64_t kNaMaxTasks = (kNaMaxBlocks + kRun - 1) / kRun;
constexpr int64_t kNaMaxStreams = kNaMaxTasks * 32;

constexpr size_t na_round(size_t b) { return (b + 255) & ~(size_t)255; }
// byte offsets inside the array (every part 256-byte aligned)
constexpr size_t kNaCtl = 0;  // int error at +0, uint32 dmgCount at +128
constexpr size_t kNaOffsets = kNaCtl + 256;
constexpr size_t kNaField = kNaOffsets + na_round(2 * sizeof(uint64_t));
constexpr size_t kNaQuant = kNaField + na_round(sizeof(DevField));
constexpr size_t kNaSeeds = kNaQuant + na_round(sizeof(DevQuant));
constexpr size_t kNaTableOf = kNaSeeds + 256;
constexpr size_t kNaJumpPow = kNaTableOf + 256;
constexpr size_t kNaJumpBytes = (size_t)kJumpPows * 128 * 2 * sizeof(uint64_t);
constexpr size_t kNaDmg = kNaJumpPow + na_round(kNaJumpBytes);
constexpr size_t kNaStates = kNaDmg + na_round((size_t)kNaMaxTasks * sizeof(uint32_t));
constexpr size_t kNaBytes = kNaStates + na_round((size_t)kNaMaxStreams * 16 + 16);
static_assert(kNaBytes == MNWCUDA_NOALLOC_SCRATCH_BYTES, "include/minnow_cuda.h states the size of the array");

// In host code this name is a stand-in, not a device address: only na_setup()'s
// cudaGetSymbolAddress result may be given to a kernel.
static __device__ __align__(256) unsigned char g_naScratch[kNaBytes];

constexpr int kNaMaxDevices = 64;
std::mutex g_naMu;
unsigned char *g_naPtr[kNaMaxDevices];  // device address of g_naScratch on device i, once set up

struct NaBegin {
    unsigned char *base;
    DevField field;
    uint64_t payloadLen;
};

// start of a call: the field record and the payload's extent arrive as kernel
// arguments (no host-to-device copy to stage), the control words start at zero
__global__ void k_na_begin(NaBegin a) {
    if (threadIdx.x != 0) return;
    *reinterpret_cast<int *>(a.base + kNaCtl) = 0;
    *reinterpret_cast<uint32_t *>(a.base + kNaCtl + 128) = 0u;
    uint64_t *off = reinterpret_cast<uint64_t *>(a.base + kNaOffsets);
    off[0] = 0;
    off[1] = a.payloadLen;
    *reinterpret_cast<DevField *>(a.base + kNaField) = a.field;
}

// end of a call: the device-side status (0 or kErrPayload) goes to the caller's word
__global__ void k_na_end(const unsigned char *base, int32_t *d_status) {
    if (threadIdx.x == 0 && d_status) *d_status = *reinterpret_cast<const int *>(base + kNaCtl);
}

// once per device, with that device current: resolve the array, tabulate J^(2^k)
int na_setup(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_naMu);
    if (g_naPtr[device]) {
        *out = g_naPtr[device];
        return 0;
    }
    unsigned char *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_naScratch);
    if (e != cudaSuccess || !ptr) {
        fprintf(stderr, "mnwcuda_DecodeFieldNoAlloc: cudaGetSymbolAddress failed on device %d (%s).\n", device,
                cudaGetErrorString(e));
        return -8;
    }
    static uint64_t pow[(size_t)kJumpPows * 256];  // as jump_powers(), into static storage
    for (int i = 0; i < 128; i++) {
        uint64_t s[2] = {i < 64 ? 1ull << i : 0ull, i >= 64 ? 1ull << (i - 64) : 0ull};
        host_jump(s);
        pow[(size_t)i * 2] = s[0];
        pow[(size_t)i * 2 + 1] = s[1];
    }
    for (int k = 1; k < kJumpPows; k++) {
        const uint64_t *prev = pow + (size_t)(k - 1) * 256;
        uint64_t *cur = pow + (size_t)k * 256;
        for (int i = 0; i < 128; i++) {
            uint64_t s[2] = {0, 0};
            for (int w = 0; w < 2; w++)
                for (int j = 0; j < 64; j++)
                    if ((prev[i * 2 + w] >> j) & 1ull) {
                        s[0] ^= prev[(w * 64 + j) * 2];
                        s[1] ^= prev[(w * 64 + j) * 2 + 1];
                    }
            cur[i * 2] = s[0];
            cur[i * 2 + 1] = s[1];
        }
    }
    e = cudaMemcpy(ptr + kNaJumpPow, pow, kNaJumpBytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        fprintf(stderr, "mnwcuda_DecodeFieldNoAlloc: set-up copy failed on device %d (%s).\n", device, cudaGetErrorString(e));
        return -8;
    }
    g_naPtr[device] = ptr;
    *out = ptr;
    return 0;
}

}  // namespace

extern "C" int mnwcuda_DecodeFieldNoAlloc(int device, const uint8_t *d_payload, int64_t payloadLen,
                                          int32_t particleLen, uint32_t fieldCode, float *d_out,
                                          int32_t *d_status, void *stream) {
    if (particleLen < 0 || payloadLen < 0 || !d_payload || (!d_out && particleLen > 0)) return -1;
    if (fieldCode != field_Posn && fieldCode != field_Velc) return -3;
    if (particleLen > kNaMaxParticles) {
        fprintf(stderr, "mnwcuda_DecodeFieldNoAlloc: ParticleLen %d is beyond the %d particles the static scratch "
                        "holds (MNWCUDA_NOALLOC_MAX_PARTICLES); use mnwcuda_DecodeFieldsDev.\n",
                particleLen, kNaMaxParticles);
        return -7;
    }
    if (((uintptr_t)d_payload & 15) || ((uintptr_t)d_out & 3)) return -6;
    if (payloadLen < mnwcuda_FieldFixedBytes(particleLen)) return -1;
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return -8;
    if (device >= kNaMaxDevices) return -1;
    if (cudaSetDevice(device) != cudaSuccess) return -8;
    unsigned char *base = nullptr;
    if (int rc = na_setup(device, &base)) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);

    NaBegin a;
    memset(&a, 0, sizeof(a));
    a.base = base;
    a.payloadLen = (uint64_t)payloadLen;
    a.field.n = particleLen;
    a.field.nb = (int32_t)nblocks_of(particleLen);
    a.field.fieldCode = fieldCode;
    const int64_t nTasks = (3 * (int64_t)a.field.nb + kRun - 1) / kRun;

    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.in = d_payload;
    p.offsets = reinterpret_cast<const uint64_t *>(base + kNaOffsets);
    p.fields = reinterpret_cast<const DevField *>(base + kNaField);
    p.nFields = 1;
    p.uniformTasks = (int)nTasks;
    p.quant = reinterpret_cast<DevQuant *>(base + kNaQuant);
    p.seeds = reinterpret_cast<uint64_t *>(base + kNaSeeds);
    p.tableOf = reinterpret_cast<int *>(base + kNaTableOf);
    p.states = reinterpret_cast<uint64_t *>(base + kNaStates);
    p.maxStreams = nTasks * 32;
    p.jumpPow = reinterpret_cast<const uint64_t *>(base + kNaJumpPow);
    p.out = d_out;
    p.error = reinterpret_cast<int *>(base + kNaCtl);
    p.dmgCount = reinterpret_cast<uint32_t *>(base + kNaCtl + 128);
    p.dmgList = reinterpret_cast<uint32_t *>(base + kNaDmg);

    k_na_begin<<<1, 32, 0, s>>>(a);
    k_parse_headers<<<1, 128, 0, s>>>(p);
    if (nTasks > 0) {
        k_stream_states<<<dim3((unsigned)((p.maxStreams + 127) / 128), 1), 128, 0, s>>>(p);
        k_unpack_dequant<false><<<(unsigned)((nTasks + kDecWarps - 1) / kDecWarps), kDecWarps * 32, 0, s>>>(p, nTasks);
        k_damaged_tasks<false><<<1, 128, 0, s>>>(p);
    }
    k_na_end<<<1, 32, 0, s>>>(base, d_status);
    return cudaGetLastError() == cudaSuccess ? 0 : -8;
}
