// Table ingest: Kubernetes quantity strings -> the integer / fp64 columns of the placement
// tables, bit-faithful to the reference's parsers:
//   ContextManager._parse_cpu            scheduler.py:172-176
//   ContextManager._parse_memory         scheduler.py:178-187
//   node row arithmetic                  scheduler.py:139-151
//   CustomScheduler._convert_pod_to_spec scheduler.py:731-764 (fp64 sums in container order;
//                                        memory strings without Ki/Mi/Gi contribute 0)
//
// Strings arrive as Arrow-style columns (one byte buffer + int64 offsets).  The device handles
// the grammar  [+-]? digits [ '.' digits ]  (at least one digit, value digits < 2^53, at most 22
// fractional digits), for which  float(text) == (double)D / 10^k  with ONE correctly rounded
// division (both operands exact) -- the same value CPython's correctly rounded float() returns.
// Anything else float() might accept or reject (exponents, underscores, blanks, inf/nan,
// non-ASCII digits, 18-digit mantissas ...) is flagged KP_ROW_HOST and left to the host mirror,
// which calls the real float().  Rows whose exact fp64 value has no integer wire value
// (e.g. a 0.1 + 0.2 cores sum) are flagged *_INEXACT; their fp64 outputs are still the
// reference's values.
#include "kp_common.cuh"

namespace kp {
namespace {

__constant__ double c_pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

constexpr unsigned long long kMaxMantissa = 1ull << 53;

// float(text[0:len]) for the device grammar; false = not in the grammar
__device__ bool parse_decimal(const uint8_t *s, int len, double *out)
{
    int i = 0;
    bool neg = false;
    if (len > 0 && (s[0] == '+' || s[0] == '-')) {
        neg = s[0] == '-';
        i = 1;
    }
    // the first nine digits cannot overflow 32 bits: one IMAD per character, no range check;
    // anything longer continues in 64 bits with the 2^53 check
    unsigned d32 = 0;
    int ndig = 0, frac = 0;
    unsigned bad = 0, points = 0;
    for (; i < len && ndig < 9; ++i) {
        const unsigned c = s[i];
        const unsigned v = c - '0';
        const unsigned isd = v <= 9u, isp = c == '.';
        bad |= (isd | isp) ^ 1u;
        points += isp;
        frac += (int)(isd & (points != 0u));
        ndig += (int)isd;
        d32 = isd ? d32 * 10u + v : d32;
    }
    unsigned long long d = d32;
    for (; i < len; ++i) {
        const unsigned c = s[i];
        if (c == '.') {
            ++points;
            continue;
        }
        const unsigned v = c - '0';
        if (v > 9u || d >= kMaxMantissa / 10) return false;  // not a digit, or the next digit could pass 2^53
        d = d * 10 + v;
        ++ndig;
        frac += points != 0u;
    }
    if (bad | (points > 1u)) return false;
    if (ndig == 0 || frac > 22) return false;
    const double mag = frac ? __ddiv_rn((double)d, c_pow10[frac]) : (double)d;
    *out = neg ? -mag : mag;
    return true;
}

__device__ __forceinline__ bool ends_with(const uint8_t *s, int len, char a, char b)
{
    return len >= 2 && s[len - 2] == (uint8_t)a && s[len - 1] == (uint8_t)b;
}

// scheduler.py:172-176 (pod form :742-745: an empty string counts as 0)
__device__ bool cpu_value(const uint8_t *s, int len, bool pod_form, double *out)
{
    if (len >= 1 && s[len - 1] == 'm') {
        double f;
        if (!parse_decimal(s, len - 1, &f)) return false;
        *out = __ddiv_rn(f, 1000.0);
        return true;
    }
    if (pod_form && len == 0) {
        *out = 0.0;
        return true;
    }
    return parse_decimal(s, len, out);
}

// scheduler.py:178-187 (pod form :747-753: no Ki/Mi/Gi suffix -> contributes nothing)
// Division by 1024 is exact for these magnitudes, so one multiplication by the power of two
// gives the same bits as the reference's chain of divisions.
__device__ bool mem_value(const uint8_t *s, int len, bool pod_form, double *out, bool *counted)
{
    *counted = true;
    double f;
    if (ends_with(s, len, 'K', 'i')) {
        if (!parse_decimal(s, len - 2, &f)) return false;
        *out = __dmul_rn(f, 9.5367431640625e-07);                  // 2^-20
        return true;
    }
    if (ends_with(s, len, 'M', 'i')) {
        if (!parse_decimal(s, len - 2, &f)) return false;
        *out = __dmul_rn(f, 0.0009765625);                         // 2^-10
        return true;
    }
    if (ends_with(s, len, 'G', 'i')) return parse_decimal(s, len - 2, out);
    if (pod_form) {
        *counted = false;
        *out = 0.0;
        return true;
    }
    if (!parse_decimal(s, len, &f)) return false;
    *out = bytes_to_gib_f64(f);
    return true;
}

// integer wire values that reproduce an fp64 value through the kernels' own conversions
__device__ bool milli_of(double cores, long long *out)
{
    if (!(fabs(cores) < 8796093022208.0)) return false;              // 2^43 cores: milli < 2^53, and
                                                                     // consecutive milli values stay distinct doubles
    const long long m0 = __double2ll_rn(__dmul_rn(cores, 1000.0));
    const long long cand[3] = {m0, m0 - 1, m0 + 1};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        if (__double_as_longlong(req_cores(cand[i])) == __double_as_longlong(cores)) {
            *out = cand[i];
            return true;
        }
    }
    return false;
}
__device__ bool bytes_of(double gib, long long *out)
{
    if (!(fabs(gib) < 8388608.0)) return false;                      // |bytes| < 2^53
    const double b = __dmul_rn(gib, 1073741824.0);                   // exact scaling
    if (b != trunc(b)) return false;
    const long long v = (long long)b;
    if (__double_as_longlong(req_gib(v)) != __double_as_longlong(gib)) return false;   // -0.0
    *out = v;
    return true;
}

struct StrCol {
    const uint8_t *text;
    const long long *off;
};

__global__ void __launch_bounds__(256)
ingest_pods_kernel(long long n_pods, const long long *container_begin, StrCol cpu, StrCol mem,
                   long long *out_milli, long long *out_bytes, double *out_cpu, double *out_mem,
                   uint8_t *out_status)
{
    for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < n_pods;
         p += (long long)gridDim.x * blockDim.x) {
        const long long c0 = container_begin[p], c1 = container_begin[p + 1];
        double cpu_sum = 0.0, mem_sum = 0.0;                         // scheduler.py:734-735
        unsigned status = 0;
        for (long long c = c0; c < c1; ++c) {
            double v;
            const long long a0 = cpu.off[c], a1 = cpu.off[c + 1];
            if (a1 < a0 || a1 - a0 > 64 || !cpu_value(cpu.text + a0, (int)(a1 - a0), true, &v)) status |= KP_ROW_HOST;
            else cpu_sum = __dadd_rn(cpu_sum, v);
            const long long b0 = mem.off[c], b1 = mem.off[c + 1];
            bool counted;
            if (b1 < b0 || b1 - b0 > 64 || !mem_value(mem.text + b0, (int)(b1 - b0), true, &v, &counted)) status |= KP_ROW_HOST;
            else if (counted) mem_sum = __dadd_rn(mem_sum, v);
        }
        long long milli = 0, bytes = 0;
        if (!(status & KP_ROW_HOST)) {
            if (!milli_of(cpu_sum, &milli)) status |= KP_ROW_CPU_INEXACT;
            if (!bytes_of(mem_sum, &bytes)) status |= KP_ROW_MEM_INEXACT;
        }
        if (out_milli) out_milli[p] = milli;
        if (out_bytes) out_bytes[p] = bytes;
        if (out_cpu) out_cpu[p] = cpu_sum;
        if (out_mem) out_mem[p] = mem_sum;
        out_status[p] = (uint8_t)status;
    }
}

// int(text) for  [+-]? digits
__device__ bool parse_int(const uint8_t *s, int len, long long *out)
{
    int i = 0;
    bool neg = false;
    if (len > 0 && (s[0] == '+' || s[0] == '-')) {
        neg = s[0] == '-';
        i = 1;
    }
    if (i >= len) return false;
    unsigned long long d = 0;
    for (; i < len; ++i) {
        const unsigned v = (unsigned)s[i] - '0';
        if (v > 9u || d >= kMaxMantissa / 10) return false;
        d = d * 10 + v;
    }
    *out = neg ? -(long long)d : (long long)d;
    return true;
}

__global__ void __launch_bounds__(256)
ingest_nodes_kernel(long long n_nodes, StrCol cpu, StrCol mem, StrCol pods, const long long *pod_count,
                    long long *out_milli, long long *out_bytes, long long *out_max_pods, double *out_cu,
                    double *out_mu, double *out_avail_cpu, double *out_avail_mem, uint8_t *out_status)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_nodes;
         i += (long long)gridDim.x * blockDim.x) {
        unsigned status = 0;
        double cores = 0.0, gib = 0.0;
        long long mp = 0, milli = 0, bytes = 0;
        bool counted;
        const long long a0 = cpu.off[i], a1 = cpu.off[i + 1];
        if (a1 < a0 || a1 - a0 > 64 || !cpu_value(cpu.text + a0, (int)(a1 - a0), false, &cores)) status |= KP_ROW_HOST;
        const long long b0 = mem.off[i], b1 = mem.off[i + 1];
        if (b1 < b0 || b1 - b0 > 64 || !mem_value(mem.text + b0, (int)(b1 - b0), false, &gib, &counted)) status |= KP_ROW_HOST;
        const long long p0 = pods.off[i], p1 = pods.off[i + 1];
        if (p1 < p0 || p1 - p0 > 32 || !parse_int(pods.text + p0, (int)(p1 - p0), &mp)) status |= KP_ROW_HOST;
        if (!(status & KP_ROW_HOST)) {
            if (!milli_of(cores, &milli)) status |= KP_ROW_CPU_INEXACT;
            if (!bytes_of(gib, &bytes)) status |= KP_ROW_MEM_INEXACT;
        }
        const double usage = (status & KP_ROW_HOST) ? 0.0 : usage_a3(pod_count[i], mp);   // scheduler.py:150-151
        if (out_milli) out_milli[i] = milli;
        if (out_bytes) out_bytes[i] = bytes;
        if (out_max_pods) out_max_pods[i] = mp;
        if (out_cu) out_cu[i] = usage;
        if (out_mu) out_mu[i] = usage;
        if (out_avail_cpu) out_avail_cpu[i] = cores;
        if (out_avail_mem) out_avail_mem[i] = gib;
        out_status[i] = (uint8_t)status;
    }
}

int grid_for(long long rows)
{
    const long long blocks = (rows + 255) / 256;
    const long long cap = (long long)sm_count_cached() * 8;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

int ingest_pods_device(const kp_pod_request_strings &in, int64_t *out_milli, int64_t *out_bytes,
                       double *out_cpu, double *out_mem, uint8_t *out_status, cudaStream_t stream)
{
    if (in.n_pods == 0) return KP_OK;
    ingest_pods_kernel<<<grid_for(in.n_pods), 256, 0, stream>>>(
        in.n_pods, (const long long *)in.container_begin,
        StrCol{in.cpu_text, (const long long *)in.cpu_offsets},
        StrCol{in.mem_text, (const long long *)in.mem_offsets}, (long long *)out_milli,
        (long long *)out_bytes, out_cpu, out_mem, out_status);
    KP_LAUNCH_CHECK("ingest_pods_kernel");
    return KP_OK;
}

int ingest_nodes_device(const kp_node_strings &in, int64_t *out_milli, int64_t *out_bytes,
                        int64_t *out_max_pods, double *out_cu, double *out_mu, double *out_avail_cpu,
                        double *out_avail_mem, uint8_t *out_status, cudaStream_t stream)
{
    if (in.n_nodes == 0) return KP_OK;
    ingest_nodes_kernel<<<grid_for(in.n_nodes), 256, 0, stream>>>(
        in.n_nodes, StrCol{in.cpu_text, (const long long *)in.cpu_offsets},
        StrCol{in.mem_text, (const long long *)in.mem_offsets},
        StrCol{in.pods_text, (const long long *)in.pods_offsets}, (const long long *)in.pod_count,
        (long long *)out_milli, (long long *)out_bytes, (long long *)out_max_pods, out_cu, out_mu,
        out_avail_cpu, out_avail_mem, out_status);
    KP_LAUNCH_CHECK("ingest_nodes_kernel");
    return KP_OK;
}

}  // namespace kp
