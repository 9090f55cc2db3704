// otedama_b200.cu — libotedama_b200.so: host runtime + C ABI over the sm_100a
// kernels (scan_kernel.cuh, job_kernels.cuh, intpeak.cuh).
//
// Host side mirrors internal/miner.Worker (internal/miner/worker.go:91-282): one
// worker per device owns a host thread, a CUDA stream and a pair of device share
// rings; SetWork swaps the job and bumps a generation, shares flow back through a
// bounded queue with a drop counter, Stats is a lock-light snapshot.  There is no
// CPU hashing path in this library: candidates are hashed, confirmed and turned
// into 32-byte hashes on the device.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/otedama_b200.h"
#include "intpeak.cuh"
#include "job_kernels.cuh"
#include "scan_kernel.cuh"
#include "scan_variants.h"
#include "scan_vs_kernel.cuh"

using namespace otd;

// ------------------------------------------------------------------ errors --
static thread_local std::string g_last_error;
static int fail(int code, const std::string& msg)
{
    g_last_error = msg;
    return code;
}
#define CU_TRY(expr)                                                                          \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail(OTD_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));       \
    } while (0)

extern "C" const char* otd_last_error(void) { return g_last_error.c_str(); }
extern "C" int otd_abi_version(void) { return OTD_ABI_VERSION; }
extern "C" int otd_abi_struct_size(int which)
{
    switch (which) {
        case 0: return (int)sizeof(otd_device_info_t);
        case 1: return (int)sizeof(otd_header_t);
        case 2: return (int)sizeof(otd_work_t);
        case 3: return (int)sizeof(otd_share_t);
        case 4: return (int)sizeof(otd_worker_config_t);
        case 5: return (int)sizeof(otd_stats_t);
        case 6: return (int)sizeof(otd_telemetry_t);
        default: return -1;
    }
}

// ----------------------------------------------------------------- devices --
extern "C" int otd_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) {
        cudaGetLastError();
        return 0;
    }
    if (e != cudaSuccess) return fail(OTD_ECUDA, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
    return n;
}

// Synchronous entry points make their device current for the duration of the call and put the
// caller's device back on return (a host application has its own idea of the current device).
struct DeviceScope {
    int prev = -1;
    DeviceScope()
    {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            prev = -1;
            cudaGetLastError();
        }
    }
    ~DeviceScope()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

static int check_device(int idx)
{
    int n = otd_device_count();
    if (n < 0) return n;
    if (n == 0) return fail(OTD_ENODEV, "no CUDA device available (this library has no CPU fallback)");
    if (idx < 0 || idx >= n) return fail(OTD_ENODEV, "device index " + std::to_string(idx) + " out of range");
    return OTD_OK;
}

extern "C" int otd_device_info(int idx, otd_device_info_t* out)
{
    if (!out) return fail(OTD_EINVAL, "otd_device_info: out is NULL");
    int rc = check_device(idx);
    if (rc) return rc;
    cudaDeviceProp p;
    CU_TRY(cudaGetDeviceProperties(&p, idx));
    memset(out, 0, sizeof *out);
    snprintf(out->id, sizeof out->id, "gpu-cuda%d", idx);
    snprintf(out->model, sizeof out->model, "%.127s", p.name);
    snprintf(out->vendor, sizeof out->vendor, "nvidia");
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, idx) == cudaSuccess) snprintf(out->pci_bus_id, sizeof out->pci_bus_id, "%s", bus);
    out->sm_count = p.multiProcessorCount;
    out->cc_major = p.major;
    out->cc_minor = p.minor;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, idx);
    out->sm_clock_khz = khz;
    out->total_mem_bytes = p.totalGlobalMem;
    // the kernels are built for sm_100a only
    out->cap_sha256d = (p.major == 10) ? 1 : 0;
    out->cap_general_compute = 1;
    return OTD_OK;
}

// --------------------------------------------------------------- telemetry --
// NVML is bound at run time so the library has no link-time dependency on it.
namespace {
struct Nvml {
    void* lib = nullptr;
    int (*init)() = nullptr;
    int (*by_bus)(const char*, void**) = nullptr;
    int (*power)(void*, unsigned int*) = nullptr;
    int (*clock)(void*, int, unsigned int*) = nullptr;
    int (*temp)(void*, int, unsigned int*) = nullptr;
    bool ok = false;
};
Nvml& nvml()
{
    static Nvml n = [] {
        Nvml m;
        m.lib = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
        if (!m.lib) return m;
        m.init = (int (*)())dlsym(m.lib, "nvmlInit_v2");
        m.by_bus = (int (*)(const char*, void**))dlsym(m.lib, "nvmlDeviceGetHandleByPciBusId_v2");
        m.power = (int (*)(void*, unsigned int*))dlsym(m.lib, "nvmlDeviceGetPowerUsage");
        m.clock = (int (*)(void*, int, unsigned int*))dlsym(m.lib, "nvmlDeviceGetClockInfo");
        m.temp = (int (*)(void*, int, unsigned int*))dlsym(m.lib, "nvmlDeviceGetTemperature");
        m.ok = m.init && m.by_bus && m.power && m.clock && m.temp && m.init() == 0;
        return m;
    }();
    return n;
}
}  // namespace

extern "C" int otd_device_telemetry(int device_index, otd_telemetry_t* out)
{
    if (!out) return fail(OTD_EINVAL, "otd_device_telemetry: out is NULL");
    int rc = check_device(device_index);
    if (rc) return rc;
    Nvml& n = nvml();
    if (!n.ok) return fail(OTD_ENODEV, "NVML (libnvidia-ml.so.1) is not available");
    char bus[32] = {0};
    CU_TRY(cudaDeviceGetPCIBusId(bus, sizeof bus, device_index));
    void* h = nullptr;
    if (n.by_bus(bus, &h) != 0 || !h) return fail(OTD_ENODEV, std::string("NVML has no device at ") + bus);
    unsigned int mw = 0, mhz = 0, t = 0;
    out->power_w = n.power(h, &mw) == 0 ? mw / 1000.0 : -1.0;
    out->sm_clock_mhz = n.clock(h, 1 /* NVML_CLOCK_SM */, &mhz) == 0 ? (double)mhz : -1.0;
    out->temperature_c = n.temp(h, 0 /* NVML_TEMPERATURE_GPU */, &t) == 0 ? (double)t : -1.0;
    return OTD_OK;
}

// ------------------------------------------------------------- header math --
static inline void put_le32(uint8_t* p, uint32_t v) { p[0] = v; p[1] = v >> 8; p[2] = v >> 16; p[3] = v >> 24; }
static inline uint32_t get_le32(const uint8_t* p)
{
    return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
}

extern "C" void otd_header_bytes(const otd_header_t* h, uint8_t out[80])
{
    put_le32(out, h->version);
    memcpy(out + 4, h->prev_hash, 32);
    memcpy(out + 36, h->merkle_root, 32);
    put_le32(out + 68, h->time);
    put_le32(out + 72, h->bits);
    put_le32(out + 76, h->nonce);
}
extern "C" void otd_parse_header(const uint8_t in[80], otd_header_t* h)
{
    h->version = get_le32(in);
    memcpy(h->prev_hash, in + 4, 32);
    memcpy(h->merkle_root, in + 36, 32);
    h->time = get_le32(in + 68);
    h->bits = get_le32(in + 72);
    h->nonce = get_le32(in + 76);
}
extern "C" int otd_hash_less_or_equal(const uint8_t h[32], const uint8_t o[32])
{
    for (int i = 31; i >= 0; i--) {
        if (h[i] < o[i]) return 1;
        if (h[i] > o[i]) return 0;
    }
    return 1;
}

// Small unsigned big integer (little-endian 32-bit limbs) for the once-per-job
// target conversions (the reference uses math/big).
struct BigU {
    static const int N = 24;  // 768 bits
    uint32_t v[N];
    BigU() { memset(v, 0, sizeof v); }
    int bit_length() const
    {
        for (int i = N - 1; i >= 0; i--)
            if (v[i]) return 32 * i + (32 - __builtin_clz(v[i]));
        return 0;
    }
    bool is_zero() const { return bit_length() == 0; }
    void shl(int s)
    {
        int ws = s / 32, bs = s % 32;
        uint32_t r[N] = {0};
        for (int i = N - 1; i >= ws; i--) {
            uint64_t x = (uint64_t)v[i - ws] << bs;
            if (bs && i - ws - 1 >= 0) x |= v[i - ws - 1] >> (32 - bs);
            r[i] = (uint32_t)x;
        }
        memcpy(v, r, sizeof v);
    }
    void shr(int s)
    {
        int ws = s / 32, bs = s % 32;
        uint32_t r[N] = {0};
        for (int i = 0; i + ws < N; i++) {
            uint64_t x = v[i + ws] >> bs;
            if (bs && i + ws + 1 < N) x |= (uint64_t)v[i + ws + 1] << (32 - bs);
            r[i] = (uint32_t)x;
        }
        memcpy(v, r, sizeof v);
    }
    bool low_bits_nonzero(int s) const  // any of the low s bits set
    {
        for (int i = 0; i < N && 32 * i < s; i++) {
            uint32_t m = (s - 32 * i >= 32) ? 0xffffffffu : ((1u << (s - 32 * i)) - 1u);
            if (v[i] & m) return true;
        }
        return false;
    }
    bool bit(int i) const { return (v[i / 32] >> (i % 32)) & 1u; }
    void add1()
    {
        for (int i = 0; i < N; i++)
            if (++v[i]) break;
    }
    uint64_t divmod_u64(uint64_t d)  // this /= d, returns remainder
    {
        unsigned __int128 rem = 0;
        for (int i = N - 1; i >= 0; i--) {
            unsigned __int128 cur = (rem << 32) | v[i];
            v[i] = (uint32_t)(cur / d);
            rem = cur % d;
        }
        return (uint64_t)rem;
    }
};

extern "C" int otd_target_from_nbits(uint32_t nbits, uint8_t target[32])
{
    char buf[96];
    const int exp = (int)(nbits >> 24);
    const uint32_t mant = nbits & 0x007fffffu;
    if (nbits & 0x00800000u) {
        snprintf(buf, sizeof buf, "miner: nBits 0x%08X has negative mantissa bit set", nbits);
        return fail(OTD_EINVAL, buf);
    }
    if (exp < 3) {
        snprintf(buf, sizeof buf, "miner: nBits 0x%08X exponent %d is below minimum 3", nbits, exp);
        return fail(OTD_EINVAL, buf);
    }
    if (mant == 0) {
        snprintf(buf, sizeof buf, "miner: nBits 0x%08X has zero mantissa", nbits);
        return fail(OTD_EINVAL, buf);
    }
    BigU b;
    b.v[0] = mant;
    const int shift = 8 * (exp - 3);
    if (shift + 23 >= 32 * BigU::N) return fail(OTD_ERANGE, "miner: nBits target overflows 256 bits");
    b.shl(shift);
    if (b.bit_length() > 256) {
        snprintf(buf, sizeof buf, "miner: nBits 0x%08X target overflows 256 bits", nbits);
        return fail(OTD_ERANGE, buf);
    }
    for (int i = 0; i < 8; i++) put_le32(target + 4 * i, b.v[i]);
    return OTD_OK;
}

extern "C" uint32_t otd_nbits_from_target(const uint8_t target[32])
{
    int len = 32;
    while (len > 0 && target[len - 1] == 0) len--;  // big-endian magnitude length
    if (len == 0) return 0;
    uint8_t b[34];
    int n = 0;
    if (target[len - 1] & 0x80) b[n++] = 0;
    for (int i = len - 1; i >= 0; i--) b[n++] = target[i];
    uint32_t mant;
    if (n == 1) mant = b[0];
    else if (n == 2) mant = (uint32_t)b[0] << 8 | b[1];
    else mant = (uint32_t)b[0] << 16 | (uint32_t)b[1] << 8 | b[2];
    return ((uint32_t)(n & 0xff) << 24) | mant;
}

extern "C" int otd_target_from_difficulty(double difficulty, uint8_t target[32])
{
    char buf[96];
    if (!(difficulty > 0) || std::isinf(difficulty)) {
        snprintf(buf, sizeof buf, "miner: invalid difficulty %g", difficulty);
        return fail(OTD_EINVAL, buf);
    }
    // difficulty = m * 2^x exactly (m < 2^53).  quotient = 0xffff * 2^(208-x) / m,
    // rounded to a 256-bit mantissa (nearest even) like a prec-256 big.Float, then
    // truncated toward zero by Float.Int.
    int e2;
    const double fr = std::frexp(difficulty, &e2);
    uint64_t m = (uint64_t)std::ldexp(fr, 53);
    int x = e2 - 53;
    while ((m & 1) == 0) { m >>= 1; x++; }
    const int S = 352;  // numerator 0xffff<<S: quotient keeps >= 299 bits
    BigU q;
    q.v[0] = 0xffffu;
    q.shl(S);
    const uint64_t rem = q.divmod_u64(m);
    const int L = q.bit_length();
    const int shift = L - 256;  // >= 2
    const bool sticky = rem != 0;
    const bool half = q.bit(shift - 1);
    const bool below = q.low_bits_nonzero(shift - 1) || sticky;
    BigU mant = q;
    mant.shr(shift);
    if (half && (below || mant.bit(0))) mant.add1();
    // value = mant * 2^(shift + 208 - x - S)
    const long exp2 = (long)shift + 208 - x - S;
    if (exp2 > 0) {
        if (mant.bit_length() + exp2 > 256) {
            snprintf(buf, sizeof buf, "miner: difficulty %g target overflows 256 bits", difficulty);
            return fail(OTD_ERANGE, buf);
        }
        mant.shl((int)exp2);
    } else if (exp2 < 0) {
        if (-exp2 >= 32 * BigU::N) mant = BigU();
        else mant.shr((int)-exp2);
    }
    if (mant.is_zero()) {
        snprintf(buf, sizeof buf, "miner: difficulty %g produces a non-positive target", difficulty);
        return fail(OTD_ERANGE, buf);
    }
    if (mant.bit_length() > 256) {
        snprintf(buf, sizeof buf, "miner: difficulty %g target overflows 256 bits", difficulty);
        return fail(OTD_ERANGE, buf);
    }
    for (int i = 0; i < 8; i++) put_le32(target + 4 * i, mant.v[i]);
    return OTD_OK;
}

extern "C" int otd_meets_target(const uint8_t hash[32], uint32_t nbits)
{
    uint8_t t[32];
    int rc = otd_target_from_nbits(nbits, t);
    if (rc) return OTD_EINVAL;
    return otd_hash_less_or_equal(hash, t);
}

// --------------------------------------------------------- job -> work glue --
extern "C" int otd_v1_job_target(uint32_t nbits, double difficulty, uint8_t target[32])
{
    int rc = otd_target_from_nbits(nbits, target);
    if (rc) return rc;
    if (difficulty > 0) {
        uint8_t dt[32];
        if (otd_target_from_difficulty(difficulty, dt) == OTD_OK) memcpy(target, dt, 32);
    }
    return OTD_OK;
}

extern "C" int otd_work_from_v2_job(uint32_t job_id, uint32_t channel_id, uint32_t version,
                                    const uint8_t merkle_root[32], const uint8_t prev_hash[32], uint32_t prev_nbits,
                                    uint32_t ntime, const uint8_t share_target[32], otd_work_t* work)
{
    if (!merkle_root || !prev_hash || !share_target || !work) return fail(OTD_EINVAL, "otd_work_from_v2_job: NULL argument");
    uint8_t target[32];
    memcpy(target, share_target, 32);
    bool zero = true;
    for (int i = 0; i < 32; i++) zero = zero && target[i] == 0;
    if (zero) {
        int rc = otd_target_from_nbits(prev_nbits, target);
        if (rc) return rc;
    }
    otd_work_t w;
    memset(&w, 0, sizeof w);
    w.job_id = job_id;
    w.channel_id = channel_id;
    w.header.version = version;
    memcpy(w.header.prev_hash, prev_hash, 32);
    memcpy(w.header.merkle_root, merkle_root, 32);
    w.header.time = ntime;
    w.header.bits = prev_nbits;
    w.nbits = prev_nbits;
    memcpy(w.target, target, 32);
    *work = w;
    return OTD_OK;
}

extern "C" int otd_work_from_v1_job(const char* job_id, uint32_t channel_id, const uint8_t merkle_root[32],
                                    uint32_t ntime, uint32_t nbits, double difficulty, otd_work_t* work)
{
    if (!job_id || !merkle_root || !work) return fail(OTD_EINVAL, "otd_work_from_v1_job: NULL argument");
    uint8_t target[32];
    int rc = otd_v1_job_target(nbits, difficulty, target);
    if (rc) return fail(rc, std::string("engine: bad target for job \"") + job_id + "\": " + otd_last_error());
    // fmt.Sscanf(job.JobID, "%d", &jobID) with a uint32 destination: optional sign, decimal digits,
    // leading spaces skipped, trailing text ignored, value must fit
    const char* p = job_id;
    while (*p == ' ' || *p == '\t') p++;
    bool neg = false;
    if (*p == '+' || *p == '-') { neg = *p == '-'; p++; }
    if (*p < '0' || *p > '9' || neg)
        return fail(OTD_EINVAL, std::string("engine: unparseable job ID \"") + job_id + "\"");
    unsigned long long v = 0;
    while (*p >= '0' && *p <= '9') {
        v = v * 10 + (unsigned)(*p - '0');
        if (v > 0xffffffffull) return fail(OTD_EINVAL, std::string("engine: unparseable job ID \"") + job_id + "\"");
        p++;
    }
    otd_work_t w;
    memset(&w, 0, sizeof w);
    w.job_id = (uint32_t)v;
    w.channel_id = channel_id;
    memcpy(w.header.merkle_root, merkle_root, 32);
    w.header.time = ntime;
    w.header.bits = nbits;
    w.nbits = nbits;
    memcpy(w.target, target, 32);
    *work = w;
    return OTD_OK;
}

static int hex_nibble(char c)
{
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    return -1;
}
// decodes the whole string; returns the byte count or -1
static long hex_decode(const char* s, uint8_t* out, size_t cap)
{
    const size_t n = strlen(s);
    if (n % 2) return -1;
    if (n / 2 > cap) return -2;
    for (size_t i = 0; i < n / 2; i++) {
        const int a = hex_nibble(s[2 * i]), b = hex_nibble(s[2 * i + 1]);
        if (a < 0 || b < 0) return -1;
        out[i] = (uint8_t)(a << 4 | b);
    }
    return (long)(n / 2);
}
static int hex_u32_be(const char* s, uint32_t* out)
{
    uint8_t b[4];
    if (!s || strlen(s) != 8 || hex_decode(s, b, 4) != 4) return -1;
    *out = (uint32_t)b[0] << 24 | (uint32_t)b[1] << 16 | (uint32_t)b[2] << 8 | b[3];
    return 0;
}

extern "C" int otd_work_from_v1_notify(const otd_v1_notify_t* n, const char* extranonce1_hex,
                                       uint32_t extranonce2_size, double difficulty, uint32_t channel_id,
                                       otd_work_t* work, uint8_t* scratch, size_t scratch_len, size_t* scratch_needed)
{
    if (!n || !work || !n->job_id || !n->prevhash_hex || !n->coinb1_hex || !n->coinb2_hex || !n->version_hex ||
        !n->nbits_hex || !n->ntime_hex || (n->n_branches && !n->merkle_branch_hex))
        return fail(OTD_EINVAL, "otd_work_from_v1_notify: NULL field");
    if (!extranonce1_hex) extranonce1_hex = "";
    if (extranonce2_size > 16) return fail(OTD_EINVAL, "otd_work_from_v1_notify: extranonce2_size > 16");
    if (n->n_branches > 64) return fail(OTD_EINVAL, "otd_work_from_v1_notify: more than 64 merkle branches");
    const size_t l1 = strlen(n->coinb1_hex) / 2, l2 = strlen(n->coinb2_hex) / 2, le1 = strlen(extranonce1_hex) / 2;
    const size_t need = l1 + le1 + l2 + 32 * (size_t)n->n_branches;
    if (scratch_needed) *scratch_needed = need;
    if (need > scratch_len || (need && !scratch))
        return fail(OTD_ERANGE, "otd_work_from_v1_notify: scratch too small (" + std::to_string(need) + " bytes needed)");
    uint32_t version, nbits, ntime;
    if (hex_u32_be(n->version_hex, &version) || hex_u32_be(n->nbits_hex, &nbits) || hex_u32_be(n->ntime_hex, &ntime))
        return fail(OTD_EINVAL, "otd_work_from_v1_notify: version / nbits / ntime must be 8 hex characters");
    uint8_t prev_wire[32];
    if (strlen(n->prevhash_hex) != 64 || hex_decode(n->prevhash_hex, prev_wire, 32) != 32)
        return fail(OTD_EINVAL, "otd_work_from_v1_notify: prevhash must be 64 hex characters");
    if (strlen(n->job_id) >= OTD_JOB_ID_MAX)
        return fail(OTD_EINVAL, "otd_work_from_v1_notify: job id longer than " + std::to_string(OTD_JOB_ID_MAX - 1) + " characters");
    otd_work_t w;
    memset(&w, 0, sizeof w);
    int rc = otd_v1_job_target(nbits, difficulty, w.target);
    if (rc) return fail(rc, std::string("engine: bad target for job \"") + n->job_id + "\": " + otd_last_error());
    // The pool's job id is an opaque string: kept verbatim for mining.submit.  The numeric id is a convenience
    // for decimal ids only (what the reference's uint32 JobID can hold); anything else leaves it 0.
    snprintf(w.job_id_str, sizeof w.job_id_str, "%s", n->job_id);
    {
        unsigned long long v = 0;
        const char* q = n->job_id;
        bool dec = *q != 0;
        for (; *q && dec; q++) {
            dec = *q >= '0' && *q <= '9';
            if (dec) v = v * 10 + (unsigned)(*q - '0');
            if (v > 0xffffffffull) dec = false;
        }
        w.job_id = dec ? (uint32_t)v : 0u;
    }
    w.channel_id = channel_id;
    w.header.time = ntime;
    w.header.bits = nbits;
    w.nbits = nbits;
    w.header.version = version;
    for (int wi = 0; wi < 8; wi++)      // each 4-byte word reversed -> header (internal) byte order
        for (int b = 0; b < 4; b++) w.header.prev_hash[4 * wi + b] = prev_wire[4 * wi + 3 - b];
    uint8_t* p = scratch;
    if (hex_decode(n->coinb1_hex, p, l1) != (long)l1) return fail(OTD_EINVAL, "otd_work_from_v1_notify: coinb1 is not hex");
    w.coinb1 = p; w.coinb1_len = (uint32_t)l1; p += l1;
    if (hex_decode(extranonce1_hex, p, le1) != (long)le1) return fail(OTD_EINVAL, "otd_work_from_v1_notify: extranonce1 is not hex");
    w.extranonce1 = p; w.extranonce1_len = (uint32_t)le1; p += le1;
    if (hex_decode(n->coinb2_hex, p, l2) != (long)l2) return fail(OTD_EINVAL, "otd_work_from_v1_notify: coinb2 is not hex");
    w.coinb2 = p; w.coinb2_len = (uint32_t)l2; p += l2;
    w.merkle_branches = p;
    w.n_branches = n->n_branches;
    for (uint32_t i = 0; i < n->n_branches; i++) {
        if (!n->merkle_branch_hex[i] || strlen(n->merkle_branch_hex[i]) != 64 || hex_decode(n->merkle_branch_hex[i], p, 32) != 32)
            return fail(OTD_EINVAL, "otd_work_from_v1_notify: merkle branch " + std::to_string(i) + " must be 64 hex characters");
        p += 32;
    }
    w.extranonce2_size = extranonce2_size;
    w.extranonce2_begin = 0;
    // every value the field can hold, capped at 2^40 values (x 2^32 nonces each: far beyond any job's life)
    w.extranonce2_count = extranonce2_size >= 5 ? (1ull << 40) : (1ull << (8 * extranonce2_size));
    w.extranonce2_stride = 1;
    *work = w;
    return OTD_OK;
}

// ------------------------------------------------------------ scan variants --
struct ScanVariant {
    const char* name;
    int v, iter;
    unsigned knobs;
    bool half;  // experiment: stops after hash 1, does not compute SHA-256d
    int block;  // threads per block
    otd_scan_launch_fn fn;
};
#define OTD_TABLE_ROW(idx, name, v, iter, kn, half, minb, block) {name, v, iter, kn, half, block, otd_scan_launch_##idx},
static const ScanVariant kAllVariants[] = {OTD_SCAN_VARIANTS(OTD_TABLE_ROW)};
#undef OTD_TABLE_ROW
static const int kNumAllVariants = (int)(sizeof kAllVariants / sizeof kAllVariants[0]);
static const int kDefaultVariant = OTD_DEFAULT_VARIANT;
static const ScanVariant* const kVariants = kAllVariants;

// Production variants are [0, OTD_NUM_PRODUCTION_VARIANTS); the experimental ones behind
// them are reachable by index only when OTD_ALLOW_EXPERIMENTAL=1 (performance studies).
static int num_usable_variants()
{
    static const int n = [] {
        const char* e = getenv("OTD_ALLOW_EXPERIMENTAL");
        return (e && e[0] == '1') ? kNumAllVariants : OTD_NUM_PRODUCTION_VARIANTS;
    }();
    return n;
}

extern "C" int otd_scan_variants(char (*names)[32], int max_names)
{
    if (names)
        for (int i = 0; i < num_usable_variants() && i < max_names; i++) snprintf(names[i], 32, "%s", kVariants[i].name);
    return num_usable_variants();
}
extern "C" int otd_scan_default_variant(void) { return kDefaultVariant; }

// Version-shared variants (scan_vs_kernel.cuh): used when a work rolls BIP-320 version bits.
struct VsVariant {
    const char* name;
    int iter, block;
    otd_vs_launch_fn fn;
};
#define OTD_VS_ROW(idx, name, iter, kn, minb, block) {name, iter, block, otd_vs_launch_##idx},
static const VsVariant kVsVariants[] = {OTD_VS_VARIANTS(OTD_VS_ROW)};
#undef OTD_VS_ROW
static const int kNumVsVariants = (int)(sizeof kVsVariants / sizeof kVsVariants[0]);
static std::atomic<int> g_vs_choice{-1};  // -1 library default, -2 off, >= 0 a variant index

extern "C" int otd_vs_variants(char (*names)[32], int max_names)
{
    if (names)
        for (int i = 0; i < kNumVsVariants && i < max_names; i++) snprintf(names[i], 32, "%s", kVsVariants[i].name);
    return kNumVsVariants;
}
extern "C" int otd_vs_default_variant(void) { return OTD_DEFAULT_VS_VARIANT; }
extern "C" int otd_vs_select(int variant)
{
    if (variant < -2 || variant >= kNumVsVariants) return fail(OTD_EINVAL, "otd_vs_select: unknown variant");
    g_vs_choice.store(variant);
    return OTD_OK;
}
// the variant a run uses for rolled work, or nullptr when sharing is switched off
static const VsVariant* vs_variant_in_use()
{
    const int c = g_vs_choice.load();
    if (c == -2) return nullptr;
    return &kVsVariants[c < 0 ? OTD_DEFAULT_VS_VARIANT : c];
}

// -------------------------------------------------------------- work copy --
struct WorkCopy {
    otd_work_t w;
    std::vector<uint8_t> blob;  // coinb1 || extranonce1 || coinb2 || branches
    bool has_coinbase = false;
    uint64_t n_en2 = 1, n_ver = 1, n_inst = 1;
    uint64_t nonce_count = 1ull << 32;
    uint32_t ver_stride = 1;
    uint64_t en2_stride = 1;
    uint32_t target_le[8];
};

static int copy_work(const otd_work_t* in, WorkCopy* wc)
{
    if (!in) return fail(OTD_EINVAL, "work is NULL");
    wc->w = *in;
    wc->w.job_id_str[OTD_JOB_ID_MAX - 1] = 0;
    const otd_work_t& w = wc->w;
    wc->nonce_count = w.nonce_count ? w.nonce_count : (1ull << 32);
    if ((uint64_t)w.nonce_begin + wc->nonce_count > (1ull << 32))
        return fail(OTD_EINVAL, "nonce window exceeds 2^32");
    for (int i = 0; i < 8; i++) wc->target_le[i] = get_le32(w.target + 4 * i);
    // version rolling
    wc->n_ver = 1;
    wc->ver_stride = w.version_stride ? w.version_stride : 1;
    if (w.version_mask) {
        const uint64_t total = 1ull << __builtin_popcount(w.version_mask);
        if (w.version_begin >= total) return fail(OTD_EINVAL, "version_begin outside the mask's index space");
        const uint64_t avail = (total - w.version_begin + wc->ver_stride - 1) / wc->ver_stride;
        wc->n_ver = w.version_count ? w.version_count : avail;
        if (wc->n_ver > avail) return fail(OTD_EINVAL, "version_count exceeds the mask's index space");
        // the preparation kernel indexes version instances with 32-bit arithmetic; BIP-320 uses 16 bits
        if (wc->n_ver > (1ull << 31)) return fail(OTD_EINVAL, "version slice larger than 2^31 indices (narrow the mask or set version_count)");
    } else if (w.version_count > 1) {
        return fail(OTD_EINVAL, "version_count > 1 without a version_mask");
    }
    // coinbase
    wc->has_coinbase = w.coinb1 != nullptr;
    wc->n_en2 = 1;
    wc->en2_stride = w.extranonce2_stride ? w.extranonce2_stride : 1;
    if (wc->has_coinbase) {
        if (w.extranonce2_size > 16) return fail(OTD_EINVAL, "extranonce2_size > 16");
        if ((w.extranonce1_len && !w.extranonce1) || (w.coinb2_len && !w.coinb2) || (w.n_branches && !w.merkle_branches))
            return fail(OTD_EINVAL, "NULL coinbase component with non-zero length");
        if ((uint64_t)w.coinb1_len + w.extranonce1_len + w.coinb2_len > (1u << 20) || w.n_branches > 64)
            return fail(OTD_EINVAL, "coinbase too large");
        if (w.extranonce2_count == 0) return fail(OTD_EINVAL, "extranonce2_count must be >= 1 with a coinbase");
        if (w.extranonce2_size < 8) {
            const uint64_t lim = 1ull << (8 * w.extranonce2_size);
            const uint64_t last = w.extranonce2_begin + (w.extranonce2_count - 1) * wc->en2_stride;
            if (w.extranonce2_begin >= lim || last >= lim || last < w.extranonce2_begin)
                return fail(OTD_EINVAL, "extranonce2 range does not fit extranonce2_size bytes");
        }
        wc->n_en2 = w.extranonce2_count;
        wc->blob.resize((size_t)w.coinb1_len + w.extranonce1_len + w.coinb2_len + 32ull * w.n_branches);
        uint8_t* p = wc->blob.data();
        memcpy(p, w.coinb1, w.coinb1_len); p += w.coinb1_len;
        if (w.extranonce1_len) memcpy(p, w.extranonce1, w.extranonce1_len);
        p += w.extranonce1_len;
        if (w.coinb2_len) memcpy(p, w.coinb2, w.coinb2_len);
        p += w.coinb2_len;
        if (w.n_branches) memcpy(p, w.merkle_branches, 32ull * w.n_branches);
        // the copies own the bytes now
        wc->w.coinb1 = wc->w.extranonce1 = wc->w.coinb2 = wc->w.merkle_branches = nullptr;
    }
    if (wc->n_en2 > (1ull << 62) / wc->n_ver) return fail(OTD_EINVAL, "search space too large");
    wc->n_inst = wc->n_en2 * wc->n_ver;
    return OTD_OK;
}

static_assert(OTD_VS_MAX_INSTANCES == (unsigned)kVsMaxVersions, "public constant out of step with the kernel's limit");
// How one run cuts a work's (instance x nonce) space into kernel launches.  Pure host arithmetic, shared by
// Engine::run and otd_plan_launches (which exposes it to CPU-side property tests).
struct LaunchGroup {
    uint32_t grp;     // version instances in the launch (1 = plain kernel)
    bool shared;      // version-shared kernel
    uint64_t step;    // nonces per launch
};
static const uint32_t kPrepBatch = 256;  // instances prepared (K3/K2) at a time
// `inst`: absolute instance index; `left_in_batch`: instances from inst to the end of its preparation batch;
// `vs_tile` / `plain_tile`: nonces per thread block of the two kernels (vs_tile 0 = sharing switched off).
static LaunchGroup plan_group(const WorkCopy& wc, uint64_t inst, uint32_t left_in_batch, uint64_t chunk,
                              uint64_t vs_tile, uint64_t plain_tile)
{
    LaunchGroup g{1, false, chunk};
    (void)plain_tile;
    // Version instances of one extranonce2 value share block 2 of hash 1: up to kVsMaxVersions of them go into
    // ONE version-shared launch (scan_vs_kernel.cuh) — never across extranonce2 values (different merkle roots)
    // nor across a preparation batch; a lone instance takes the plain kernel.
    if (vs_tile && wc.w.version_mask)
        g.grp = (uint32_t)std::min<uint64_t>({(uint64_t)kVsMaxVersions, wc.n_ver - inst % wc.n_ver, (uint64_t)left_in_batch});
    g.shared = g.grp >= 2;
    // nonces per launch: the configured candidate count split over the group, in whole blocks
    if (g.shared) g.step = std::max(vs_tile, chunk / g.grp / vs_tile * vs_tile);
    return g;
}

extern "C" int otd_work_partition(const otd_work_t* in, uint32_t g, uint32_t G, otd_work_t* out)
{
    if (!in || !out || G == 0 || g >= G) return fail(OTD_EINVAL, "otd_work_partition: bad argument");
    WorkCopy wc;  // validates the input and resolves defaults (counts, strides)
    int rc = copy_work(in, &wc);
    if (rc) return rc;
    otd_work_t w = *in;
    if (in->version_mask) {
        if (wc.n_ver < G) return fail(OTD_EINVAL, "otd_work_partition: fewer version indices than devices");
        w.version_begin = in->version_begin + g * wc.ver_stride;
        w.version_stride = wc.ver_stride * G;
        w.version_count = (uint32_t)((wc.n_ver - g + G - 1) / G);
    } else if (in->coinb1) {
        if (wc.n_en2 < G) return fail(OTD_EINVAL, "otd_work_partition: fewer extranonce2 values than devices");
        w.extranonce2_begin = in->extranonce2_begin + (uint64_t)g * wc.en2_stride;
        w.extranonce2_stride = wc.en2_stride * G;
        w.extranonce2_count = (wc.n_en2 - g + G - 1) / G;
    } else {
        if (wc.nonce_count < G) return fail(OTD_EINVAL, "otd_work_partition: fewer nonces than devices");
        const uint64_t lo = wc.nonce_count * g / G, hi = wc.nonce_count * (g + 1) / G;
        w.nonce_begin = in->nonce_begin + (uint32_t)lo;
        w.nonce_count = hi - lo;
    }
    *out = w;
    return OTD_OK;
}

extern "C" int otd_plan_launches(const otd_work_t* work, uint32_t launch_log2, otd_launch_plan_t* out,
                                 uint32_t max_launches, uint64_t* n_launches)
{
    WorkCopy wc;
    int rc = copy_work(work, &wc);
    if (rc) return rc;
    if (max_launches && !out) return fail(OTD_EINVAL, "otd_plan_launches: out is NULL");
    if (n_launches && wc.n_inst > (1ull << 22))
        return fail(OTD_ERANGE, "otd_plan_launches: counting every launch of more than 2^22 instances is refused "
                                "(pass n_launches = NULL for the first max_launches entries, or bound the work)");
    if (launch_log2 == 0) launch_log2 = 32;
    launch_log2 = std::min(32u, std::max(16u, launch_log2));
    const uint64_t chunk = 1ull << launch_log2;
    const VsVariant* const vs = (wc.w.flags & OTD_WORK_PER_VERSION_LAUNCH) ? nullptr : vs_variant_in_use();
    const ScanVariant& kv = kVariants[kDefaultVariant];
    uint64_t n = 0;
    uint32_t written = 0;
    for (uint64_t bf = 0; bf < wc.n_inst; bf += kPrepBatch) {
        const uint32_t nb = (uint32_t)std::min<uint64_t>(kPrepBatch, wc.n_inst - bf);
        for (uint32_t i = 0; i < nb;) {
            const LaunchGroup lg = plan_group(wc, bf + i, nb - i, chunk, vs ? (uint64_t)vs->block * vs->iter : 0,
                                              (uint64_t)kv.block * kv.v * kv.iter);
            for (uint64_t off = 0; off < wc.nonce_count; off += lg.step) {
                if (written < max_launches) {
                    otd_launch_plan_t& p = out[written++];
                    p.first_instance = bf + i;
                    p.n_instances = lg.grp;
                    p.version_shared = lg.shared ? 1 : 0;
                    p.nonce_begin = wc.w.nonce_begin + (uint32_t)off;
                    p.nonce_count = std::min(lg.step, wc.nonce_count - off);
                    p.extranonce2 = wc.has_coinbase ? wc.w.extranonce2_begin + ((bf + i) / wc.n_ver) * wc.en2_stride : 0;
                    p.version_index_first = wc.w.version_begin + (uint32_t)((bf + i) % wc.n_ver) * wc.ver_stride;
                }
                n++;
            }
            i += lg.grp;
        }
        if (written >= max_launches && !n_launches) break;  // nothing more to report
    }
    if (n_launches) *n_launches = n;
    return (int)written;
}

extern "C" int otd_share_to_v1_submit(const otd_share_t* share, uint32_t extranonce2_size, uint32_t version_mask,
                                      otd_v1_submit_t* out)
{
    if (!share || !out) return fail(OTD_EINVAL, "otd_share_to_v1_submit: NULL argument");
    if (extranonce2_size > 16) return fail(OTD_EINVAL, "otd_share_to_v1_submit: extranonce2_size exceeds 16");
    memset(out, 0, sizeof *out);
    for (uint32_t i = 0; i < extranonce2_size; i++) {
        const unsigned b = i < 8 ? (unsigned)(share->extranonce2 >> (8 * i)) & 0xffu : 0u;  // same bytes as the coinbase got
        snprintf(out->extranonce2_hex + 2 * i, 3, "%02x", b);
    }
    snprintf(out->ntime_hex, sizeof out->ntime_hex, "%08x", share->ntime);
    snprintf(out->nonce_hex, sizeof out->nonce_hex, "%08x", share->nonce);
    if (version_mask) snprintf(out->version_bits_hex, sizeof out->version_bits_hex, "%08x", share->version & version_mask);
    if (share->job_id_str[0] && memchr(share->job_id_str, 0, sizeof share->job_id_str))
        snprintf(out->job_id, sizeof out->job_id, "%s", share->job_id_str);
    else
        snprintf(out->job_id, sizeof out->job_id, "%u", share->job_id);
    return OTD_OK;
}

// ------------------------------------------------------------------ engine --
struct RunStats {
    uint64_t hashes = 0, launches = 0, candidates = 0, found = 0, ring_dropped = 0, shared_launches = 0;
    double kernel_ms = 0, span_ms = 0;
};

struct ShareSink {
    virtual void on_share(const otd_share_t& s) = 0;
    virtual void on_launch_done(uint64_t hashes, uint64_t candidates, double ms, bool version_shared) = 0;
    virtual ~ShareSink() {}
};

struct Slot {
    RingHeader* d_hdr = nullptr;
    RingEntry* d_ent = nullptr;
    RingHeader* h_hdr = nullptr;
    RingEntry* h_ent = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr, edone = nullptr;
    bool busy = false;
    uint64_t count = 0, en2 = 0, tile = 0;  // candidates of the launch, extranonce2 value, candidates per block
    uint32_t inst0 = 0;                     // instance index (low 32 bits) of the launch's first version instance
    uint32_t n_ver = 1;                     // version instances in the launch (1 for the plain kernel)
    uint32_t versions[kVsMaxVersions] = {0};
};

class Engine {
   public:
    static const uint32_t kBatch = kPrepBatch;
    int dev = -1;
    uint32_t cap = 0;
    cudaStream_t stream = nullptr, aux = nullptr;
    Slot slot[2];
    unsigned int* d_stop = nullptr;
    unsigned int* h_gen = nullptr;  // pinned: [0] = run generation, [1] = abort value (0)
    unsigned int gen = 0;
    uint8_t* d_blob = nullptr;
    size_t blob_cap = 0;
    uint8_t* d_roots = nullptr;
    JobConsts* d_consts = nullptr;
    JobConsts* h_consts = nullptr;
    uint32_t* d_versions = nullptr;
    uint32_t* h_versions = nullptr;
    cudaEvent_t ev_first = nullptr, ev_last = nullptr;
    bool ready = false;

    int init(int device, uint32_t ring_cap)
    {
        int rc = check_device(device);
        if (rc) return rc;
        dev = device;
        cap = ring_cap ? ring_cap : 4096;
        CU_TRY(cudaSetDevice(dev));
        CU_TRY(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        CU_TRY(cudaStreamCreateWithFlags(&aux, cudaStreamNonBlocking));
        for (auto& s : slot) {
            CU_TRY(cudaMalloc(&s.d_hdr, sizeof(RingHeader)));
            CU_TRY(cudaMalloc(&s.d_ent, sizeof(RingEntry) * (size_t)cap));
            CU_TRY(cudaMallocHost(&s.h_hdr, sizeof(RingHeader)));
            CU_TRY(cudaMallocHost(&s.h_ent, sizeof(RingEntry) * (size_t)cap));
            CU_TRY(cudaEventCreate(&s.e0));
            CU_TRY(cudaEventCreate(&s.e1));
            CU_TRY(cudaEventCreate(&s.edone));
        }
        CU_TRY(cudaMalloc(&d_stop, sizeof(unsigned int)));
        CU_TRY(cudaMallocHost(&h_gen, 2 * sizeof(unsigned int)));
        h_gen[0] = 0; h_gen[1] = 0;
        CU_TRY(cudaMalloc(&d_roots, 32 * (size_t)kBatch));
        CU_TRY(cudaMalloc(&d_consts, sizeof(JobConsts) * kBatch));
        CU_TRY(cudaMallocHost(&h_consts, sizeof(JobConsts) * kBatch));
        CU_TRY(cudaMalloc(&d_versions, sizeof(uint32_t) * kBatch));
        CU_TRY(cudaMallocHost(&h_versions, sizeof(uint32_t) * kBatch));
        CU_TRY(cudaEventCreate(&ev_first));
        CU_TRY(cudaEventCreate(&ev_last));
        ready = true;
        return OTD_OK;
    }

    void destroy()
    {
        if (dev < 0) return;
        cudaSetDevice(dev);
        if (stream) cudaStreamSynchronize(stream);
        if (aux) cudaStreamSynchronize(aux);
        for (auto& s : slot) {
            cudaFree(s.d_hdr); cudaFree(s.d_ent); cudaFreeHost(s.h_hdr); cudaFreeHost(s.h_ent);
            if (s.e0) cudaEventDestroy(s.e0);
            if (s.e1) cudaEventDestroy(s.e1);
            if (s.edone) cudaEventDestroy(s.edone);
            s = Slot();
        }
        cudaFree(d_stop); cudaFreeHost(h_gen); cudaFree(d_blob); cudaFree(d_roots); cudaFree(d_consts);
        cudaFreeHost(h_consts); cudaFree(d_versions); cudaFreeHost(h_versions);
        if (ev_first) cudaEventDestroy(ev_first);
        if (ev_last) cudaEventDestroy(ev_last);
        if (stream) cudaStreamDestroy(stream);
        if (aux) cudaStreamDestroy(aux);
        stream = aux = nullptr;
        d_stop = nullptr; h_gen = nullptr; d_blob = nullptr; d_roots = nullptr; d_consts = nullptr;
        h_consts = nullptr; d_versions = nullptr; h_versions = nullptr; ev_first = ev_last = nullptr;
        dev = -1;
        ready = false;
    }

    int upload_blob(const WorkCopy& wc)
    {
        if (!wc.has_coinbase) return OTD_OK;
        const size_t need = std::max<size_t>(wc.blob.size(), 1);
        if (need > blob_cap) {
            if (d_blob) cudaFree(d_blob);
            d_blob = nullptr;
            blob_cap = 0;
            CU_TRY(cudaMalloc(&d_blob, need));
            blob_cap = need;
        }
        // synchronous small copy: the blob lives in pageable host memory
        CU_TRY(cudaMemcpyAsync(d_blob, wc.blob.data(), wc.blob.size(), cudaMemcpyHostToDevice, stream));
        CU_TRY(cudaStreamSynchronize(stream));
        return OTD_OK;
    }

    HeaderTemplate make_template(const WorkCopy& wc) const
    {
        HeaderTemplate ht;
        ht.version = wc.w.header.version;
        ht.version_mask = wc.w.version_mask;
        memcpy(ht.prev_hash, wc.w.header.prev_hash, 32);
        memcpy(ht.merkle_root, wc.w.header.merkle_root, 32);
        ht.time = wc.w.header.time;
        ht.bits = wc.w.header.bits;
        memcpy(ht.target, wc.target_le, sizeof ht.target);
        return ht;
    }

    // Fill h_consts / h_versions for instances [first, first+n). Leaves the stream idle.
    int prepare_batch(const WorkCopy& wc, uint64_t first, uint32_t n, uint8_t* roots_out /*host, may be NULL*/)
    {
        const bool rolled = wc.has_coinbase || wc.w.version_mask;
        if (!rolled) {
            // plain header: midstate and folding once on the host (north_star: "header
            // midstate ... precomputed once per job on the host")
            uint8_t hdr[80];
            otd_header_bytes(&wc.w.header, hdr);
            uint32_t mid[8], tail[3];
            header_midstate(hdr, mid, tail);
            fold_job_consts(mid, tail[0], tail[1], tail[2], wc.target_le, &h_consts[0]);
            h_versions[0] = wc.w.header.version;
            if (roots_out) memcpy(roots_out, wc.w.header.merkle_root, 32);
            return OTD_OK;
        }
        const uint32_t n_ver = (uint32_t)wc.n_ver;
        const uint64_t e_first = first / wc.n_ver;
        const uint64_t e_last = (first + n - 1) / wc.n_ver;
        const uint32_t n_e = (uint32_t)(e_last - e_first + 1);  // <= n <= kBatch
        const uint8_t* roots = nullptr;
        if (wc.has_coinbase) {
            CoinbaseDesc cd;
            cd.blob = d_blob;
            cd.l1 = wc.w.coinb1_len; cd.le1 = wc.w.extranonce1_len; cd.le2 = wc.w.extranonce2_size;
            cd.l2 = wc.w.coinb2_len; cd.n_branches = wc.w.n_branches;
            roll_extranonce2_kernel<<<(n_e + 63) / 64, 64, 0, stream>>>(
                cd, wc.w.extranonce2_begin + e_first * wc.en2_stride, wc.en2_stride, n_e, d_roots);
            CU_TRY(cudaGetLastError());
            roots = d_roots;
        }
        instance_consts_kernel<<<(n + 63) / 64, 64, 0, stream>>>(make_template(wc), roots, (uint32_t)e_first, n_ver,
                                                                   wc.w.version_begin, wc.ver_stride, first, n,
                                                                   d_consts, d_versions);
        CU_TRY(cudaGetLastError());
        CU_TRY(cudaMemcpyAsync(h_consts, d_consts, sizeof(JobConsts) * n, cudaMemcpyDeviceToHost, stream));
        CU_TRY(cudaMemcpyAsync(h_versions, d_versions, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, stream));
        CU_TRY(cudaStreamSynchronize(stream));
        if (roots_out) {
            if (wc.has_coinbase) {
                std::vector<uint8_t> tmp(32 * (size_t)n_e);
                CU_TRY(cudaMemcpy(tmp.data(), d_roots, tmp.size(), cudaMemcpyDeviceToHost));
                for (uint32_t i = 0; i < n; i++)
                    memcpy(roots_out + 32 * (size_t)i, tmp.data() + 32 * (size_t)((first + i) / wc.n_ver - e_first), 32);
            } else {
                for (uint32_t i = 0; i < n; i++) memcpy(roots_out + 32 * (size_t)i, wc.w.header.merkle_root, 32);
            }
        }
        return OTD_OK;
    }

    int harvest(Slot& s, const WorkCopy& wc, uint64_t seq, ShareSink& sink, RunStats& rs,
                const std::function<bool()>& aborted, bool& abort_sent)
    {
        // poll instead of blocking so an abort request reaches the device promptly
        for (;;) {
            cudaError_t q = cudaEventQuery(s.edone);
            if (q == cudaSuccess) break;
            if (q != cudaErrorNotReady) return fail(OTD_ECUDA, std::string("scan kernel: ") + cudaGetErrorString(q));
            if (!abort_sent && aborted()) {
                CU_TRY(cudaMemcpyAsync(d_stop, &h_gen[1], sizeof(unsigned int), cudaMemcpyHostToDevice, aux));
                abort_sent = true;
            }
            std::this_thread::sleep_for(std::chrono::microseconds(100));
        }
        float ms = 0;
        CU_TRY(cudaEventElapsedTime(&ms, s.e0, s.e1));
        const uint32_t pushed = s.h_hdr->count;
        const uint32_t n = std::min(pushed, cap);
        if (n) {
            CU_TRY(cudaMemcpyAsync(s.h_ent, s.d_ent, sizeof(RingEntry) * (size_t)n, cudaMemcpyDeviceToHost, aux));
            CU_TRY(cudaStreamSynchronize(aux));
        }
        const uint64_t hashed = std::min<uint64_t>((uint64_t)s.h_hdr->blocks_done * s.tile, s.count);
        rs.hashes += hashed;
        rs.launches += 1;
        rs.candidates += s.h_hdr->candidates;
        rs.kernel_ms += ms;
        rs.found += pushed;
        rs.ring_dropped += pushed - n;
        sink.on_launch_done(hashed, s.h_hdr->candidates, ms, s.n_ver >= 2);
        for (uint32_t i = 0; i < n; i++) {
            otd_share_t sh;
            memset(&sh, 0, sizeof sh);
            sh.channel_id = wc.w.channel_id;
            sh.job_id = wc.w.job_id;
            memcpy(sh.job_id_str, wc.w.job_id_str, sizeof sh.job_id_str);
            sh.nonce = s.h_ent[i].nonce;
            sh.ntime = wc.w.header.time;
            const uint32_t vi = s.h_ent[i].instance - s.inst0;  // which of the launch's version instances
            if (vi >= s.n_ver) return fail(OTD_ECUDA, "scan kernel: ring entry names an instance outside its launch");
            sh.version = s.versions[vi];
            memcpy(sh.hash, s.h_ent[i].hash, 32);
            sh.extranonce2 = s.en2;
            sh.work_seq = seq;
            sh.device_index = dev;
            sink.on_share(sh);
        }
        s.busy = false;
        return OTD_OK;
    }

    // One pass over the work's whole search space.  Returns OTD_OK also when aborted.
    int run(const WorkCopy& wc, uint64_t seq, int variant, uint32_t launch_log2, ShareSink& sink,
            const std::function<bool()>& aborted, RunStats& rs)
    {
        if (!ready) return fail(OTD_EINVAL, "engine not initialised");
        if (variant < 0 || variant >= num_usable_variants()) variant = kDefaultVariant;
        const ScanVariant& kv = kVariants[variant];
        const VsVariant* const vs = (wc.w.flags & OTD_WORK_PER_VERSION_LAUNCH) ? nullptr : vs_variant_in_use();
        if (launch_log2 == 0) launch_log2 = 32;
        launch_log2 = std::min(32u, std::max(16u, launch_log2));
        const uint64_t chunk = 1ull << launch_log2;
        CU_TRY(cudaSetDevice(dev));
        int rc = upload_blob(wc);
        if (rc) return rc;
        // A previous run's abort write travels on `aux`: make sure it has landed before this run's
        // generation is written, or it could overwrite it and make every block of this run exit.
        CU_TRY(cudaStreamSynchronize(aux));
        CU_TRY(cudaStreamSynchronize(stream));
        for (auto& s : slot) s.busy = false;  // nothing of an earlier (possibly failed) run is in flight now
        gen += 1;
        if (gen == 0) gen = 1;
        h_gen[0] = gen;
        CU_TRY(cudaMemcpyAsync(d_stop, &h_gen[0], sizeof(unsigned int), cudaMemcpyHostToDevice, stream));
        CU_TRY(cudaStreamSynchronize(stream));
        bool abort_sent = false, first_launch = true, halted = false;
        int next = 0;
        for (uint64_t bf = 0; bf < wc.n_inst && !abort_sent && !halted; bf += kBatch) {
            const uint32_t nb = (uint32_t)std::min<uint64_t>(kBatch, wc.n_inst - bf);
            for (auto& s : slot)
                if (s.busy && (rc = harvest(s, wc, seq, sink, rs, aborted, abort_sent))) return rc;
            if (abort_sent || aborted()) break;
            if ((rc = prepare_batch(wc, bf, nb, nullptr))) return rc;
            for (uint32_t i = 0; i < nb && !abort_sent && !halted;) {
                const uint64_t inst = bf + i;
                const uint64_t en2 = wc.w.extranonce2_begin + (inst / wc.n_ver) * wc.en2_stride;
                const LaunchGroup lg = plan_group(wc, inst, nb - i, chunk, vs ? (uint64_t)vs->block * vs->iter : 0,
                                                  (uint64_t)kv.block * kv.v * kv.iter);
                const uint32_t grp = lg.grp;
                const bool shared = lg.shared;
                const uint64_t step = lg.step;
                const uint64_t blk_nonces = shared ? (uint64_t)vs->block * vs->iter : (uint64_t)kv.block * kv.v * kv.iter;
                VsParams vp;
                if (shared) memcpy(vp.jc, &h_consts[i], sizeof(JobConsts) * grp);
                for (uint64_t off = 0; off < wc.nonce_count; off += step) {
                    Slot& s = slot[next];
                    next ^= 1;
                    if (s.busy && (rc = harvest(s, wc, seq, sink, rs, aborted, abort_sent))) return rc;
                    if (abort_sent || aborted()) { halted = true; break; }
                    const uint64_t cnt = std::min(step, wc.nonce_count - off);
                    CU_TRY(cudaMemsetAsync(s.d_hdr, 0, sizeof(RingHeader), stream));
                    if (first_launch) { CU_TRY(cudaEventRecord(ev_first, stream)); first_launch = false; }
                    CU_TRY(cudaEventRecord(s.e0, stream));
                    cudaError_t le = shared
                        ? vs->fn(vp, grp, wc.w.nonce_begin + (uint32_t)off, cnt, (uint32_t)(inst & 0xffffffffu), s.d_hdr,
                                 s.d_ent, cap, d_stop, gen, stream)
                        : kv.fn(h_consts[i], wc.w.nonce_begin + (uint32_t)off, cnt, (uint32_t)(inst & 0xffffffffu),
                                s.d_hdr, s.d_ent, cap, d_stop, gen, stream);
                    if (le != cudaSuccess) return fail(OTD_ECUDA, std::string("scan launch: ") + cudaGetErrorString(le));
                    CU_TRY(cudaEventRecord(s.e1, stream));
                    CU_TRY(cudaEventRecord(ev_last, stream));
                    CU_TRY(cudaMemcpyAsync(s.h_hdr, s.d_hdr, sizeof(RingHeader), cudaMemcpyDeviceToHost, stream));
                    CU_TRY(cudaEventRecord(s.edone, stream));
                    s.busy = true;
                    s.count = cnt * grp;
                    s.tile = blk_nonces * grp;
                    s.en2 = wc.has_coinbase ? en2 : 0;
                    s.inst0 = (uint32_t)(inst & 0xffffffffu);
                    s.n_ver = grp;
                    for (uint32_t k = 0; k < grp; k++) s.versions[k] = h_versions[i + k];
                    if (shared) rs.shared_launches += 1;
                }
                i += grp;
            }
        }
        for (int k = 0; k < 2; k++) {
            Slot& s = slot[next];
            next ^= 1;
            if (s.busy && (rc = harvest(s, wc, seq, sink, rs, aborted, abort_sent))) return rc;
        }
        if (!first_launch) {
            float span = 0;
            CU_TRY(cudaEventSynchronize(ev_last));
            CU_TRY(cudaEventElapsedTime(&span, ev_first, ev_last));
            rs.span_ms += span;
        }
        if (abort_sent) CU_TRY(cudaStreamSynchronize(aux));
        return OTD_OK;
    }
};

// --------------------------------------------------------- synchronous scan --
struct CollectSink : ShareSink {
    std::vector<otd_share_t> shares;
    void on_share(const otd_share_t& s) override { shares.push_back(s); }
    void on_launch_done(uint64_t, uint64_t, double, bool) override {}
};

static std::mutex g_pool_mu;
static std::vector<std::unique_ptr<Engine>> g_pool;       // one cached engine per device
static std::vector<std::unique_ptr<std::mutex>> g_pool_locks;

static int pool_engine(int dev, Engine** out, std::mutex** lock)
{
    int rc = check_device(dev);
    if (rc) return rc;
    std::lock_guard<std::mutex> g(g_pool_mu);
    if ((int)g_pool.size() <= dev) {
        g_pool.resize(dev + 1);
        while ((int)g_pool_locks.size() <= dev) g_pool_locks.emplace_back(new std::mutex());
    }
    if (!g_pool[dev]) {
        std::unique_ptr<Engine> e(new Engine());
        rc = e->init(dev, 1u << 16);
        if (rc) {
            e->destroy();  // release whatever init got before it failed
            return rc;
        }
        g_pool[dev] = std::move(e);
    }
    *out = g_pool[dev].get();
    *lock = g_pool_locks[dev].get();
    return OTD_OK;
}

extern "C" int otd_scan(int device_index, const otd_work_t* work, int kernel_variant, otd_share_t* out,
                        uint32_t max_shares, uint64_t* n_found, uint64_t* hashes, double* device_ms,
                        uint64_t* launches)
{
    DeviceScope scope;  // the caller's current device is restored on return
    WorkCopy wc;
    int rc = copy_work(work, &wc);
    if (rc) return rc;
    if (max_shares && !out) return fail(OTD_EINVAL, "otd_scan: out is NULL");
    Engine* eng;
    std::mutex* lock;
    if ((rc = pool_engine(device_index, &eng, &lock))) return rc;
    std::lock_guard<std::mutex> g(*lock);
    CollectSink sink;
    RunStats rs;
    rc = eng->run(wc, 1, kernel_variant, 32, sink, [] { return false; }, rs);
    if (rc) return rc;
    std::sort(sink.shares.begin(), sink.shares.end(), [](const otd_share_t& a, const otd_share_t& b) {
        if (a.extranonce2 != b.extranonce2) return a.extranonce2 < b.extranonce2;
        if (a.version != b.version) return a.version < b.version;
        return a.nonce < b.nonce;
    });
    const size_t n = std::min<size_t>(sink.shares.size(), max_shares);
    for (size_t i = 0; i < n; i++) out[i] = sink.shares[i];
    if (n_found) *n_found = rs.found;
    if (hashes) *hashes = rs.hashes;
    if (device_ms) *device_ms = rs.span_ms;
    if (launches) *launches = rs.launches;
    return (int)n;
}

// Burst scan of n independent headers (BASELINE configs[1] shape): all job
// constants are folded up front, then every launch is queued back to back on
// the stream between ONE pair of CUDA events — no host synchronisation inside
// the timed region, inputs resident (kernel parameters) before it starts.
extern "C" int otd_scan_headers(int device_index, const otd_header_t* headers, uint32_t n_headers,
                                const uint8_t target[32], uint32_t nonce_begin, uint64_t nonce_count,
                                int kernel_variant, otd_share_t* out, uint32_t max_shares, uint64_t* n_found,
                                uint64_t* hashes, double* device_ms, uint64_t* launches)
{
    DeviceScope scope;  // the caller's current device is restored on return
    if (!headers || !target || n_headers == 0) return fail(OTD_EINVAL, "otd_scan_headers: bad argument");
    if (max_shares && !out) return fail(OTD_EINVAL, "otd_scan_headers: out is NULL");
    if (nonce_count == 0) nonce_count = 1ull << 32;
    if ((uint64_t)nonce_begin + nonce_count > (1ull << 32)) return fail(OTD_EINVAL, "nonce window exceeds 2^32");
    Engine* eng;
    std::mutex* lock;
    int rc = pool_engine(device_index, &eng, &lock);
    if (rc) return rc;
    std::lock_guard<std::mutex> g(*lock);
    if (kernel_variant < 0 || kernel_variant >= num_usable_variants()) kernel_variant = kDefaultVariant;
    const ScanVariant& kv = kVariants[kernel_variant];
    uint32_t tgt[8];
    for (int i = 0; i < 8; i++) tgt[i] = get_le32(target + 4 * i);
    std::vector<JobConsts> jcs(n_headers);
    for (uint32_t i = 0; i < n_headers; i++) {
        uint8_t hdr[80];
        uint32_t mid[8], tail[3];
        otd_header_bytes(&headers[i], hdr);
        header_midstate(hdr, mid, tail);
        fold_job_consts(mid, tail[0], tail[1], tail[2], tgt, &jcs[i]);
    }
    CU_TRY(cudaSetDevice(device_index));
    Slot& s = eng->slot[0];
    CU_TRY(cudaStreamSynchronize(eng->aux));
    eng->gen += 1;
    if (eng->gen == 0) eng->gen = 1;
    eng->h_gen[0] = eng->gen;
    CU_TRY(cudaMemcpyAsync(eng->d_stop, &eng->h_gen[0], sizeof(unsigned int), cudaMemcpyHostToDevice, eng->stream));
    CU_TRY(cudaMemsetAsync(s.d_hdr, 0, sizeof(RingHeader), eng->stream));
    CU_TRY(cudaStreamSynchronize(eng->stream));
    CU_TRY(cudaEventRecord(eng->ev_first, eng->stream));
    for (uint32_t i = 0; i < n_headers; i++) {
        cudaError_t le = kv.fn(jcs[i], nonce_begin, nonce_count, i, s.d_hdr, s.d_ent, eng->cap, eng->d_stop, eng->gen,
                               eng->stream);
        if (le != cudaSuccess) return fail(OTD_ECUDA, std::string("scan launch: ") + cudaGetErrorString(le));
    }
    CU_TRY(cudaEventRecord(eng->ev_last, eng->stream));
    CU_TRY(cudaMemcpyAsync(s.h_hdr, s.d_hdr, sizeof(RingHeader), cudaMemcpyDeviceToHost, eng->stream));
    CU_TRY(cudaStreamSynchronize(eng->stream));
    float ms = 0;
    CU_TRY(cudaEventElapsedTime(&ms, eng->ev_first, eng->ev_last));
    const uint32_t pushed = s.h_hdr->count;
    const uint32_t n = std::min(pushed, eng->cap);
    if (n) CU_TRY(cudaMemcpy(s.h_ent, s.d_ent, sizeof(RingEntry) * (size_t)n, cudaMemcpyDeviceToHost));
    std::vector<otd_share_t> shares(n);
    for (uint32_t i = 0; i < n; i++) {
        otd_share_t& sh = shares[i];
        memset(&sh, 0, sizeof sh);
        const uint32_t hi = s.h_ent[i].instance;
        sh.job_id = hi;  // the header's index in the batch
        sh.nonce = s.h_ent[i].nonce;
        sh.ntime = headers[hi].time;
        sh.version = headers[hi].version;
        memcpy(sh.hash, s.h_ent[i].hash, 32);
        sh.work_seq = 1;
        sh.device_index = device_index;
    }
    std::sort(shares.begin(), shares.end(), [](const otd_share_t& a, const otd_share_t& b) {
        if (a.job_id != b.job_id) return a.job_id < b.job_id;
        return a.nonce < b.nonce;
    });
    const size_t nw = std::min<size_t>(shares.size(), max_shares);
    for (size_t i = 0; i < nw; i++) out[i] = shares[i];
    if (n_found) *n_found = pushed;
    if (hashes) *hashes = nonce_count * n_headers;
    if (device_ms) *device_ms = ms;
    if (launches) *launches = n_headers;
    return (int)nw;
}

extern "C" int otd_prepare_instances(int device_index, const otd_work_t* work, uint32_t first, uint32_t n,
                                     uint8_t* merkle_roots, uint32_t* versions, uint64_t* extranonce2s,
                                     uint8_t* midstates)
{
    DeviceScope scope;  // the caller's current device is restored on return
    WorkCopy wc;
    int rc = copy_work(work, &wc);
    if (rc) return rc;
    if ((uint64_t)first + n > wc.n_inst) return fail(OTD_EINVAL, "instance range outside the work's search space");
    Engine* eng;
    std::mutex* lock;
    if ((rc = pool_engine(device_index, &eng, &lock))) return rc;
    std::lock_guard<std::mutex> g(*lock);
    CU_TRY(cudaSetDevice(device_index));
    if ((rc = eng->upload_blob(wc))) return rc;
    for (uint32_t done = 0; done < n; done += Engine::kBatch) {
        const uint32_t nb = std::min<uint32_t>(Engine::kBatch, n - done);
        rc = eng->prepare_batch(wc, (uint64_t)first + done, nb, merkle_roots ? merkle_roots + 32ull * done : nullptr);
        if (rc) return rc;
        for (uint32_t i = 0; i < nb; i++) {
            const uint64_t inst = (uint64_t)first + done + i;
            if (versions) versions[done + i] = eng->h_versions[i];
            if (extranonce2s) extranonce2s[done + i] = wc.has_coinbase ? wc.w.extranonce2_begin + (inst / wc.n_ver) * wc.en2_stride : 0;
            if (midstates)
                for (int k = 0; k < 8; k++) {
                    const uint32_t v = eng->h_consts[i].mid[k];
                    uint8_t* p = midstates + 32ull * (done + i) + 4 * k;
                    p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v;
                }
        }
    }
    return OTD_OK;
}

// ------------------------------------------------------------ hash_headers --
// Batch HashHeader.  The caller's buffers are ordinary (pageable) host memory.  The call walks them in chunks
// over two streams with two device buffer pairs that are allocated once per device and kept (no
// cudaMalloc/cudaFree per call): the driver's own staged copy of chunk i+1 overlaps chunk i's kernel and
// copy-out.  The kernel itself runs at billions of headers per second; end to end the entry point is bound by
// the host-side copies of pageable memory (an extra hop through library-owned pinned staging was
// measured and is slower than letting the driver stage: 82 vs 98+ M headers/s).
namespace {
struct HashStage {
    static const size_t kChunk = 1u << 18;  // headers per chunk: 20 MiB in, 8 MiB out
    uint8_t *d_in[2] = {nullptr, nullptr}, *d_out[2] = {nullptr, nullptr};
    cudaStream_t st[2] = {nullptr, nullptr};
    bool ready = false;
    std::mutex mu;
    int init()
    {
        if (ready) return OTD_OK;
        for (int k = 0; k < 2; k++) {
            CU_TRY(cudaMalloc(&d_in[k], 80 * kChunk));
            CU_TRY(cudaMalloc(&d_out[k], 32 * kChunk));
            CU_TRY(cudaStreamCreateWithFlags(&st[k], cudaStreamNonBlocking));
        }
        ready = true;
        return OTD_OK;
    }
};
std::mutex g_hash_mu;
std::vector<std::unique_ptr<HashStage>> g_hash_stage;  // one per device, kept until exit
}  // namespace

extern "C" int otd_hash_headers(int device_index, const uint8_t* headers80, size_t n, uint8_t* hashes32)
{
    DeviceScope scope;  // the caller's current device is restored on return
    int rc = check_device(device_index);
    if (rc) return rc;
    if (n == 0) return OTD_OK;
    if (!headers80 || !hashes32) return fail(OTD_EINVAL, "otd_hash_headers: NULL buffer");
    CU_TRY(cudaSetDevice(device_index));
    HashStage* hs;
    {
        std::lock_guard<std::mutex> g(g_hash_mu);
        if ((int)g_hash_stage.size() <= device_index) g_hash_stage.resize(device_index + 1);
        if (!g_hash_stage[device_index]) g_hash_stage[device_index].reset(new HashStage());
        hs = g_hash_stage[device_index].get();
    }
    std::lock_guard<std::mutex> g(hs->mu);  // one call at a time per device: the device buffers are shared
    if ((rc = hs->init())) return rc;
    const size_t nchunks = (n + HashStage::kChunk - 1) / HashStage::kChunk;
    for (size_t c = 0; c < nchunks; c++) {
        const int k = (int)(c & 1);
        const size_t off = c * HashStage::kChunk, cnt = std::min(HashStage::kChunk, n - off);
        // stream order protects the slot: this chunk's copy-in queues behind the copy-out of chunk c-2
        CU_TRY(cudaMemcpyAsync(hs->d_in[k], headers80 + 80 * off, 80 * cnt, cudaMemcpyHostToDevice, hs->st[k]));
        hash_headers_kernel<<<(unsigned)((cnt + 127) / 128), 128, 0, hs->st[k]>>>(hs->d_in[k], cnt, hs->d_out[k]);
        CU_TRY(cudaGetLastError());
        CU_TRY(cudaMemcpyAsync(hashes32 + 32 * off, hs->d_out[k], 32 * cnt, cudaMemcpyDeviceToHost, hs->st[k]));
    }
    CU_TRY(cudaStreamSynchronize(hs->st[0]));
    CU_TRY(cudaStreamSynchronize(hs->st[1]));
    return OTD_OK;
}

// ------------------------------------------------------------------ worker --
struct otd_worker : ShareSink {
    int dev = -1;
    otd_worker_config_t cfg;
    Engine eng;
    std::thread th;
    std::mutex mu;
    std::condition_variable cv_work, cv_share;
    std::shared_ptr<WorkCopy> work;
    std::atomic<uint64_t> seq{0};
    std::atomic<bool> stop{false};
    bool exhausted = false;
    uint64_t exhausted_reported = 0;  // work_seq whose end (exhausted or failed) a poll has already returned for
    std::deque<otd_share_t> queue;
    std::atomic<uint64_t> hashes{0}, shares{0}, dropped{0}, launches{0}, candidates{0}, shared_launches{0};
    double device_ms = 0;
    std::chrono::steady_clock::time_point t_start;
    std::string thread_error;

    void on_share(const otd_share_t& s) override
    {
        shares.fetch_add(1);
        std::lock_guard<std::mutex> g(mu);
        if (queue.size() >= cfg.queue_capacity) {
            dropped.fetch_add(1);  // non-blocking send, like worker.go:271-275
            return;
        }
        queue.push_back(s);
        cv_share.notify_all();
    }
    void on_launch_done(uint64_t h, uint64_t c, double ms, bool version_shared) override
    {
        hashes.fetch_add(h);
        launches.fetch_add(1);
        if (version_shared) shared_launches.fetch_add(1);
        candidates.fetch_add(c);
        std::lock_guard<std::mutex> g(mu);
        device_ms += ms;
    }

    void main_loop()
    {
        cudaSetDevice(dev);
        for (;;) {
            std::shared_ptr<WorkCopy> wc;
            uint64_t my_seq;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_work.wait(lk, [&] { return stop.load() || (work && !exhausted); });
                if (stop.load()) return;
                wc = work;
                my_seq = seq.load();
            }
            RunStats rs;
            int rc = eng.run(*wc, my_seq, cfg.kernel_variant, cfg.launch_log2, *this,
                             [&] { return stop.load() || seq.load() != my_seq; }, rs);
            shares.fetch_add(rs.ring_dropped);  // found on the device, lost to a full ring
            dropped.fetch_add(rs.ring_dropped);
            std::lock_guard<std::mutex> g(mu);
            if (rc) {
                thread_error = otd_last_error();
                exhausted = true;  // surfaces as a zero hash rate, like a wedged device
                cv_share.notify_all();
                continue;
            }
            if (!stop.load() && seq.load() == my_seq && !(wc->w.flags & OTD_WORK_REPEAT)) {
                exhausted = true;
                cv_share.notify_all();
            }
        }
    }
};

extern "C" otd_worker_t* otd_worker_create(int device_index, const otd_worker_config_t* cfg)
{
    DeviceScope scope;  // the caller's current device is restored on return
    std::unique_ptr<otd_worker> w(new otd_worker());
    w->dev = device_index;
    otd_worker_config_t c;
    memset(&c, 0, sizeof c);
    c.kernel_variant = -1;
    if (cfg) c = *cfg;
    if (c.ring_capacity == 0) c.ring_capacity = 4096;
    if (c.queue_capacity == 0) c.queue_capacity = 65536;
    if (c.launch_log2 == 0) c.launch_log2 = 32;
    if (c.launch_log2 < 16 || c.launch_log2 > 32) {
        fail(OTD_EINVAL, "launch_log2 must be in [16, 32]");
        return nullptr;
    }
    w->cfg = c;
    if (w->eng.init(device_index, c.ring_capacity) != OTD_OK) {
        w->eng.destroy();
        return nullptr;
    }
    w->t_start = std::chrono::steady_clock::now();
    otd_worker* raw = w.release();
    raw->th = std::thread([raw] { raw->main_loop(); });
    return raw;
}

extern "C" int64_t otd_worker_set_work(otd_worker_t* w, const otd_work_t* work)
{
    if (!w) return fail(OTD_EINVAL, "worker is NULL");
    std::shared_ptr<WorkCopy> wc;
    if (work) {
        wc = std::make_shared<WorkCopy>();
        int rc = copy_work(work, wc.get());
        if (rc) return rc;
    }
    std::lock_guard<std::mutex> g(w->mu);
    w->work = wc;
    w->exhausted = false;
    w->thread_error.clear();  // a failure belongs to the work it happened on
    const uint64_t s = w->seq.fetch_add(1) + 1;
    w->cv_work.notify_all();
    w->cv_share.notify_all();
    return (int64_t)s;
}

extern "C" int otd_worker_poll_shares(otd_worker_t* w, otd_share_t* out, int max_shares, int timeout_ms)
{
    if (!w || !out || max_shares <= 0) return fail(OTD_EINVAL, "otd_worker_poll_shares: bad argument");
    std::unique_lock<std::mutex> lk(w->mu);
    // Wake for: a share, shutdown, or the END of the current work (exhausted / failed) that no poll has
    // returned for yet.  Once that end has been reported, an empty poll honours its timeout again instead
    // of returning at once for as long as the worker stays idle.
    auto ended_unreported = [&] { return w->exhausted && w->exhausted_reported != w->seq.load(); };
    auto ready = [&] { return !w->queue.empty() || ended_unreported() || w->stop.load(); };
    if (timeout_ms < 0) w->cv_share.wait(lk, ready);
    else if (timeout_ms > 0) w->cv_share.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready);
    int n = 0;
    while (n < max_shares && !w->queue.empty()) {
        out[n++] = w->queue.front();
        w->queue.pop_front();
    }
    if (n == 0 && ended_unreported()) {
        w->exhausted_reported = w->seq.load();
        if (!w->thread_error.empty()) return fail(OTD_ECUDA, w->thread_error);  // reported once per failed work
    }
    return n;
}

extern "C" int otd_worker_wait_exhausted(otd_worker_t* w, int timeout_ms)
{
    if (!w) return fail(OTD_EINVAL, "worker is NULL");
    std::unique_lock<std::mutex> lk(w->mu);
    auto done = [&] { return w->exhausted || w->stop.load(); };
    if (timeout_ms < 0) w->cv_share.wait(lk, done);
    else w->cv_share.wait_for(lk, std::chrono::milliseconds(timeout_ms), done);
    if (!w->thread_error.empty()) return fail(OTD_ECUDA, w->thread_error);
    return w->exhausted ? 1 : 0;
}

extern "C" int otd_worker_stats(otd_worker_t* w, otd_stats_t* out)
{
    if (!w || !out) return fail(OTD_EINVAL, "otd_worker_stats: bad argument");
    memset(out, 0, sizeof *out);
    const auto now = std::chrono::steady_clock::now();
    out->uptime_ns = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(now - w->t_start).count();
    out->hashes_total = w->hashes.load();
    out->shares_found = w->shares.load();
    out->shares_dropped = w->dropped.load();
    out->launches = w->launches.load();
    out->shared_launches = w->shared_launches.load();
    out->candidates = w->candidates.load();
    out->hash_rate = out->uptime_ns ? (double)out->hashes_total / ((double)out->uptime_ns * 1e-9) : 0.0;
    out->work_seq = w->seq.load();
    std::lock_guard<std::mutex> g(w->mu);
    out->device_ms = w->device_ms;
    out->has_work = w->work ? 1 : 0;
    out->exhausted = w->exhausted ? 1 : 0;
    return OTD_OK;
}

extern "C" int otd_worker_has_work(otd_worker_t* w)
{
    if (!w) return fail(OTD_EINVAL, "worker is NULL");
    std::lock_guard<std::mutex> g(w->mu);
    return w->work ? 1 : 0;
}
extern "C" int otd_worker_device_index(otd_worker_t* w) { return w ? w->dev : fail(OTD_EINVAL, "worker is NULL"); }

extern "C" int otd_worker_destroy(otd_worker_t* w)
{
    DeviceScope scope;  // the caller's current device is restored on return
    if (!w) return OTD_OK;
    {
        std::lock_guard<std::mutex> g(w->mu);
        w->stop.store(true);
        w->cv_work.notify_all();
        w->cv_share.notify_all();
    }
    if (w->th.joinable()) w->th.join();
    w->eng.destroy();
    delete w;
    return OTD_OK;
}

// ---------------------------------------------------------------- int peak --
static const char* kIntPeakNames[IPM_COUNT] = {"lop3", "shf", "iadd3", "imad", "mix_alu",
                                               "mix_dual", "mix_dual1", "imad_wide", "prmt", "alu_fma", "imad_hi",
                                               "alu_wide", "alu2_wide", "alu2_fma", "alu_wide_fma",
                                               "alu2_fma_imm", "alu2_fma_const", "alu2_fma_pair", "alu2_hi", "alu4_rot"};

extern "C" int otd_int_peak_modes(char (*names)[32], int max_names)
{
    if (names)
        for (int i = 0; i < IPM_COUNT && i < max_names; i++) snprintf(names[i], 32, "%s", kIntPeakNames[i]);
    return IPM_COUNT;
}

template <int MODE>
static cudaError_t launch_peak(int grid, int iters, uint32_t* sink, long long* cyc, cudaStream_t st)
{
    int_peak_kernel<MODE><<<grid, 1024, 0, st>>>(iters, 1u, 12345u, sink, cyc);
    return cudaGetLastError();
}

extern "C" int otd_int_peak(int device_index, int mode, int iters, double* ops_per_s, double* sm_mhz, double* ms_out)
{
    DeviceScope scope;  // the caller's current device is restored on return
    int rc = check_device(device_index);
    if (rc) return rc;
    if (mode < 0 || mode >= IPM_COUNT) return fail(OTD_EINVAL, "otd_int_peak: unknown mode");
    if (iters <= 0) iters = 2048;
    CU_TRY(cudaSetDevice(device_index));
    int sms = 0;
    CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_index));
    // one 1024-thread block per SM (8 warps per scheduler): block 0 then lives as long
    // as the kernel, so its clock64 span over the event time is the SM clock
    const int grid = sms;
    uint32_t* d_sink = nullptr;
    long long* d_cyc = nullptr;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e = cudaMalloc(&d_sink, sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&d_cyc, sizeof(long long));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&e0);
    if (e == cudaSuccess) e = cudaEventCreate(&e1);
    typedef cudaError_t (*fn_t)(int, int, uint32_t*, long long*, cudaStream_t);
    static const fn_t fns[IPM_COUNT] = {launch_peak<0>, launch_peak<1>, launch_peak<2>, launch_peak<3>, launch_peak<4>,
                                        launch_peak<5>, launch_peak<6>, launch_peak<7>, launch_peak<8>, launch_peak<9>, launch_peak<10>,
                                        launch_peak<11>, launch_peak<12>, launch_peak<13>, launch_peak<14>,
                                        launch_peak<15>, launch_peak<16>, launch_peak<17>, launch_peak<18>, launch_peak<19>};
    for (int w = 0; w < 2 && e == cudaSuccess; w++) e = fns[mode](grid, iters / 4 + 1, d_sink, d_cyc, st);  // warm-up
    float ms = 0;
    long long cyc = 0;
    if (e == cudaSuccess) e = cudaEventRecord(e0, st);
    if (e == cudaSuccess) e = fns[mode](grid, iters, d_sink, d_cyc, st);
    if (e == cudaSuccess) e = cudaEventRecord(e1, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    if (e == cudaSuccess) e = cudaMemcpy(&cyc, d_cyc, sizeof cyc, cudaMemcpyDeviceToHost);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
    cudaFree(d_sink);
    cudaFree(d_cyc);
    if (e != cudaSuccess) return fail(OTD_ECUDA, std::string("otd_int_peak: ") + cudaGetErrorString(e));
    const double ops = (double)grid * 1024.0 * (double)iters * kIpUnroll * kIpChains * ip_ops_per_step(mode);
    if (ops_per_s) *ops_per_s = ops / (ms * 1e-3);
    if (sm_mhz) *sm_mhz = (double)cyc / (ms * 1e-3) / 1e6;
    if (ms_out) *ms_out = ms;
    return OTD_OK;
}

// ----------------------------------------------------- scan without allocation --
// otd_scan_noalloc: the same search as otd_scan for a plain-header work, for
// callers that are not allowed to (or cannot) allocate device memory — a process
// whose GPU memory is handed out up front by another allocator or capped by its
// operator, or code on a latency-critical path, where cudaMalloc/cudaFree would
// synchronise the whole device or fail under memory pressure.  Nothing is
// allocated here, on the device or the host heap: the kernel's scratch (share
// ring, its counters, the stop word) lives in ONE file-scope device array that
// every GPU loading this library holds until the process exits.
//
// Layout of the array: [0,16) ring header (count / candidates / blocks_done),
// [64,68) stop word, [256, 256 + 40*OTD_NOALLOC_MAX_SHARES) ring entries.
static __device__ __align__(256) unsigned char g_noalloc_scratch[OTD_NOALLOC_SCRATCH_BYTES];

static_assert(sizeof(RingEntry) == 40 && sizeof(RingHeader) == 16, "scratch layout");
static_assert(256 + sizeof(RingEntry) * OTD_NOALLOC_MAX_SHARES <= OTD_NOALLOC_SCRATCH_BYTES, "scratch size");

static const int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mu;                               // one call at a time: every call shares the array
static unsigned char* g_noalloc_ptr[kNoallocMaxDevices];      // device address of the array, per device
static bool g_noalloc_ready[kNoallocMaxDevices];

// Set-up, once per device, with that device current.  In host code the name
// g_noalloc_scratch is only a handle for the symbol, not a device address: the
// kernel must be given what cudaGetSymbolAddress returns.
static int noalloc_setup(int dev, unsigned char** out)
{
    if (dev < 0 || dev >= kNoallocMaxDevices) return fail(OTD_EINVAL, "otd_scan_noalloc: device index out of range");
    if (!g_noalloc_ready[dev]) {
        unsigned char* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(OTD_ECUDA, std::string("otd_scan_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_noalloc_ptr[dev] = p;
        g_noalloc_ready[dev] = true;
    }
    *out = g_noalloc_ptr[dev];
    return OTD_OK;
}

extern "C" int otd_scan_noalloc(int device_index, const otd_work_t* work, int kernel_variant, otd_share_t* out,
                                uint32_t max_shares, uint64_t* n_found, uint64_t* hashes)
{
    DeviceScope scope;  // the caller's current device is restored on return
    if (!work) return fail(OTD_EINVAL, "otd_scan_noalloc: work is NULL");
    if (max_shares && !out) return fail(OTD_EINVAL, "otd_scan_noalloc: out is NULL");
    if (max_shares > OTD_NOALLOC_MAX_SHARES)
        return fail(OTD_ERANGE, "otd_scan_noalloc: max_shares " + std::to_string(max_shares) +
                                    " exceeds the fixed ring of " + std::to_string(OTD_NOALLOC_MAX_SHARES) +
                                    " entries (use otd_scan for larger result sets)");
    if (work->version_mask || work->coinb1 || work->version_count > 1)
        return fail(OTD_EINVAL, "otd_scan_noalloc: only plain-header work (no version rolling, no coinbase) is supported");
    const uint64_t nonce_count = work->nonce_count ? work->nonce_count : (1ull << 32);
    if ((uint64_t)work->nonce_begin + nonce_count > (1ull << 32))
        return fail(OTD_EINVAL, "nonce window exceeds 2^32");
    int rc = check_device(device_index);
    if (rc) return rc;
    if (kernel_variant < 0 || kernel_variant >= num_usable_variants()) kernel_variant = kDefaultVariant;
    const ScanVariant& kv = kVariants[kernel_variant];

    std::lock_guard<std::mutex> g(g_noalloc_mu);
    CU_TRY(cudaSetDevice(device_index));
    unsigned char* base = nullptr;
    if ((rc = noalloc_setup(device_index, &base))) return rc;
    RingHeader* d_hdr = reinterpret_cast<RingHeader*>(base);
    unsigned int* d_stop = reinterpret_cast<unsigned int*>(base + 64);
    RingEntry* d_ent = reinterpret_cast<RingEntry*>(base + 256);
    const uint32_t cap = max_shares;  // entries beyond the caller's buffer would be dropped anyway

    // job constants on the host stack (midstate + folding), as for any plain header
    uint8_t hdr[80];
    uint32_t mid[8], tail[3], tgt[8];
    otd_header_bytes(&work->header, hdr);
    header_midstate(hdr, mid, tail);
    for (int i = 0; i < 8; i++) tgt[i] = get_le32(work->target + 4 * i);
    JobConsts jc;
    fold_job_consts(mid, tail[0], tail[1], tail[2], tgt, &jc);

    // The array keeps the previous call's contents: the counters and the stop word must
    // start at zero (generation 0 = "keep running"); ring entries need no clearing, only
    // the first `count` of them are ever read back.
    CU_TRY(cudaMemset(base, 0, 256));
    cudaError_t le = kv.fn(jc, work->nonce_begin, nonce_count, 0u, d_hdr, d_ent, cap, d_stop, 0u, (cudaStream_t)0);
    if (le != cudaSuccess) return fail(OTD_ECUDA, std::string("otd_scan_noalloc: launch: ") + cudaGetErrorString(le));
    RingHeader h;
    CU_TRY(cudaMemcpy(&h, d_hdr, sizeof h, cudaMemcpyDeviceToHost));  // waits for the kernel
    const uint32_t n = std::min(h.count, cap);
    RingEntry chunk[128];
    for (uint32_t done = 0; done < n; done += 128) {
        const uint32_t m = std::min<uint32_t>(128, n - done);
        CU_TRY(cudaMemcpy(chunk, d_ent + done, sizeof(RingEntry) * m, cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < m; i++) {
            otd_share_t& sh = out[done + i];
            memset(&sh, 0, sizeof sh);
            sh.channel_id = work->channel_id;
            sh.job_id = work->job_id;
            memcpy(sh.job_id_str, work->job_id_str, sizeof sh.job_id_str - 1);
            sh.nonce = chunk[i].nonce;
            sh.ntime = work->header.time;
            sh.version = work->header.version;
            memcpy(sh.hash, chunk[i].hash, 32);
            sh.work_seq = 1;
            sh.device_index = device_index;
        }
    }
    std::sort(out, out + n, [](const otd_share_t& a, const otd_share_t& b) { return a.nonce < b.nonce; });
    if (n_found) *n_found = h.count;
    if (hashes) *hashes = nonce_count;
    return (int)n;
}
