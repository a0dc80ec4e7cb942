// Setup on the device: the fixed-width HITRAN records
// (parse_HITRAN_file.c:224-413; columns: molecule 0-1, isotopologue 2, v 3-14, S 15-24,
// gamma_air 35-39, gamma_self 40-44, E'' 45-54, n 55-58, delta 59-66) converted to numbers
// by one thread per record.
//
// The reference converts every field with strtod (utilities.c:149-170), i.e. correctly
// rounded.  Here a field is read as an integer mantissa m < 2^53 and a decimal exponent e and
// converted as
//     0 <= e <= 22      fl(m * 5^e) * 2^e        (both factors exact: one IEEE rounding)
//     -22 <= e < 0      fl(m / 5^-e) * 2^e       (the same)
//     -45 <= e < -22    m / (hi + lo) * 2^e      with 5^-e = hi + lo exactly (pow5_table.inc):
//                       quotient, exact remainder (fma), one correction term, one final
//                       rounding -- correct unless the true quotient lies within 2^-30 ulp of
//                       a rounding boundary, which is detected
// and anything else -- more digits, other exponents, text strtod would treat specially
// (inf, nan, hex), a boundary case, any malformed field -- is NOT decided here: the row is
// marked kRowHost and the host library converts it with strtod itself (host/hitran.c), so
// error messages and odd inputs behave exactly as before.
#include "common.cuh"

#include <cstdint>

namespace grt {

enum { kRowOk = 0, kRowBadMolecule = 1, kRowBadField = 2, kRowHost = 3 };   // host/hitran.c

struct Pow5 { double hi, lo; };
__constant__ Pow5 c_pow5[46] = {
#include "pow5_table.inc"
};

__device__ __forceinline__ bool is_space(char const c)
{
    return c == ' ' || (c >= '\t' && c <= '\r');
}

__device__ __forceinline__ bool is_digit(char const c)
{
    return c >= '0' && c <= '9';
}

__device__ __forceinline__ double pow2(int const k)
{
    return __longlong_as_double((long long)(1023 + k) << 52);
}

// m*10^e for 0 < m < 2^53, -45 <= e <= 22; false if the result cannot be guaranteed.
__device__ bool scale_decimal(uint64_t const m, int const e, double *out)
{
    double const md = (double)(long long)m;              // exact
    if (e >= 0)
    {
        *out = __dmul_rn(__dmul_rn(md, c_pow5[e].hi), pow2(e));
        return true;
    }
    int const k = -e;
    double const H = c_pow5[k].hi;
    double const L = c_pow5[k].lo;
    double const q0 = __ddiv_rn(md, H);
    if (k <= 22)
    {
        *out = __dmul_rn(q0, pow2(e));
        return true;
    }
    double const r = fma(-q0, H, md);                    // exact remainder
    double const p = __dmul_rn(q0, L);
    double const pe = fma(q0, L, -p);                    // p + pe = q0*L exactly
    double const q1 = __ddiv_rn(__dsub_rn(__dsub_rn(r, p), pe), H);
    double const x = __dadd_rn(q0, q1);
    // what the last addition rounded away, against half a unit in the last place of x
    double const t = fabs(__dadd_rn(__dsub_rn(q0, x), q1));
    long long const xb = __double_as_longlong(x);
    double const ulp = __longlong_as_double((xb & 0x7ff0000000000000ll) - (52ll << 52));
    if (fabs(__dsub_rn(t, 0.5*ulp)) < ulp*9.313225746154785e-10)   // 2^-30
    {
        return false;
    }
    *out = __dmul_rn(x, pow2(e));
    return true;
}

// strtod on s[0, width) for the plain decimal forms; false = leave the field to the host.
__device__ bool parse_decimal(char const *s, int const width, double *out)
{
    int i = 0;
    while (i < width && is_space(s[i])) ++i;
    bool neg = false;
    if (i < width && (s[i] == '+' || s[i] == '-'))
    {
        neg = s[i] == '-';
        ++i;
    }
    uint64_t m = 0;
    int e10 = 0;
    bool any = false;
    bool point = false;
    for (; i<width; ++i)
    {
        char const c = s[i];
        if (is_digit(c))
        {
            any = true;
            if (m >= 100000000000000000ull) return false;     // more digits than 2^53 holds
            m = m*10u + (uint64_t)(c - '0');
            if (point) --e10;
        }
        else if (c == '.' && !point)
        {
            point = true;
        }
        else
        {
            break;
        }
    }
    if (!any) return false;
    if (i < width && (s[i] == 'e' || s[i] == 'E'))
    {
        int j = i + 1;
        bool eneg = false;
        if (j < width && (s[j] == '+' || s[j] == '-'))
        {
            eneg = s[j] == '-';
            ++j;
        }
        if (j >= width || !is_digit(s[j])) return false;
        int ex = 0;
        for (; j<width && is_digit(s[j]); ++j)
        {
            ex = ex < 10000 ? ex*10 + (s[j] - '0') : ex;
        }
        e10 += eneg ? -ex : ex;
        i = j;
    }
    // whatever follows the number must be blank (strtod would stop there too, but text such
    // as "0x1p3" or "1.234-100" is the host's to judge)
    for (; i<width; ++i)
    {
        if (s[i] == '\0') break;
        if (!is_space(s[i])) return false;
    }
    double v = 0.;
    if (m != 0)
    {
        if (m >= (1ull << 53) || e10 > 22 || e10 < -45) return false;
        if (!scale_decimal(m, e10, &v)) return false;
    }
    *out = neg ? -v : v;
    return true;
}

// (fp_t)(float)d with the reference's range check (parse_HITRAN_file.c:150-170).
__device__ __forceinline__ bool to_float_field(double const d, real_t *out)
{
    if (!(d >= -3.4028234663852886e38 && d <= 3.4028234663852886e38)) return false;
    *out = (real_t)__double2float_rn(d);
    return true;
}

struct HitranColumns
{
    short *mol;
    int *iso;
    real_t *vnn, *snn, *yair, *yself, *en, *n, *d;
    signed char *status;
};

constexpr int kRecordHead = 67;       // the columns the library reads

__global__ void __launch_bounds__(128)
hitran_parse_kernel(char const *text, uint64_t const *start, uint64_t const count,
                    HitranColumns const out)
{
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= count) return;
    char rec[kRecordHead + 1];
    char const *src = text + start[i];
#pragma unroll 1
    for (int k=0; k<kRecordHead; ++k) rec[k] = src[k];
    rec[kRecordHead] = '\0';
    int status = kRowOk;
    // molecule id: strtol over both characters (host/hitran.c parse_rows)
    int mol = -1;
    if (is_digit(rec[1]) && (is_digit(rec[0]) || is_space(rec[0])))
    {
        mol = (is_digit(rec[0]) ? 10*(rec[0] - '0') : 0) + (rec[1] - '0');
    }
    else
    {
        status = kRowHost;
    }
    // isotopologue code (parse_HITRAN_file.c:177-194)
    int iso = 0;
    char const ci = rec[2];
    if (ci == '0') iso = 10;
    else if (ci >= 'A' && ci <= 'Z') iso = ci - 'A' + 11;
    else if (ci >= '1' && ci <= '9') iso = ci - '0';
    else status = kRowHost;
    double dv = 0., ds = 0., f = 0.;
    real_t yair = 0, yself = 0, en = 0, n = 0, d = 0;
    if (status == kRowOk)
    {
        bool ok = parse_decimal(rec + 3, 12, &dv) && parse_decimal(rec + 15, 10, &ds);
        ok = ok && parse_decimal(rec + 35, 5, &f) && to_float_field(f, &yair);
        ok = ok && parse_decimal(rec + 40, 5, &f) && to_float_field(f, &yself);
        ok = ok && parse_decimal(rec + 45, 10, &f) && to_float_field(f, &en);
        ok = ok && parse_decimal(rec + 55, 4, &f) && to_float_field(f, &n);
        ok = ok && parse_decimal(rec + 59, 8, &f) && to_float_field(f, &d);
        if (!ok) status = kRowHost;
    }
    out.status[i] = (signed char)status;
    out.mol[i] = (short)mol;
    out.iso[i] = iso;
    out.vnn[i] = (real_t)dv;
    out.snn[i] = (real_t)ds;
    out.yair[i] = yair;
    out.yself[i] = yself;
    out.en[i] = en;
    out.n[i] = n;
    out.d[i] = d;
}

// Test hook: `count` fields of `width` characters each, back to back.
__global__ void __launch_bounds__(128)
decimal_fields_kernel(char const *text, int const width, uint64_t const count, double *out,
                      signed char *status)
{
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= count) return;
    char buf[32];
    for (int k=0; k<width; ++k) buf[k] = text[i*(uint64_t)width + (uint64_t)k];
    double v = 0.;
    bool const ok = parse_decimal(buf, width, &v);
    out[i] = v;
    status[i] = ok ? kRowOk : kRowHost;
}

template <class T>
static int to_host(T *dst, T const *src, uint64_t n, cudaStream_t st)
{
    GRT_CUDA(cudaMemcpyAsync(dst, src, sizeof(T)*n, cudaMemcpyDeviceToHost, st));
    return 0;
}

}  // namespace grt

using namespace grt;

extern "C" {

int grt_hitran_convert(int device, char const *text, uint64_t size, uint64_t const *start,
                       uint64_t num_records, short *mol, int *iso, grt_real *vnn, grt_real *snn,
                       grt_real *yair, grt_real *yself, grt_real *en, grt_real *n, grt_real *d,
                       signed char *status)
{
    if (text == nullptr || start == nullptr || mol == nullptr || iso == nullptr ||
        vnn == nullptr || snn == nullptr || yair == nullptr || yself == nullptr ||
        en == nullptr || n == nullptr || d == nullptr || status == nullptr)
    {
        return fail("null argument");
    }
    if (num_records == 0) return 0;
    for (uint64_t k=0; k<num_records; k+=(num_records > 64 ? num_records/64 : 1))
    {
        if (start[k] + kRecordHead > size) return fail("record %llu lies beyond the text",
                                                      (unsigned long long)k);
    }
    if (start[num_records - 1] + kRecordHead > size) return fail("the last record is truncated");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    GRT_CUDA(cudaSetDevice(device));
    uint64_t const c = num_records;
    // one allocation: text | start | the ten columns (8-byte aligned pieces)
    auto pad = [](uint64_t b) { return (b + 255u) & ~(uint64_t)255u; };
    uint64_t const o_start = pad(size);
    uint64_t const o_real = o_start + pad(8*c);
    uint64_t const o_iso = o_real + 7*pad(sizeof(real_t)*c);
    uint64_t const o_mol = o_iso + pad(4*c);
    uint64_t const o_status = o_mol + pad(2*c);
    uint64_t const total = o_status + pad(c);
    char *buf = nullptr;
    GRT_CUDA(cudaMalloc(&buf, total));
    int rc = 0;
    do
    {
        cudaError_t e = cudaMemcpyAsync(buf, text, size, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(buf + o_start, start, 8*c, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { rc = fail("cuda: %s (upload of the HITRAN text)", cudaGetErrorString(e)); break; }
        HitranColumns col;
        real_t *r = reinterpret_cast<real_t *>(buf + o_real);
        uint64_t const rs = pad(sizeof(real_t)*c)/sizeof(real_t);
        col.vnn = r; col.snn = r + rs; col.yair = r + 2*rs; col.yself = r + 3*rs;
        col.en = r + 4*rs; col.n = r + 5*rs; col.d = r + 6*rs;
        col.iso = reinterpret_cast<int *>(buf + o_iso);
        col.mol = reinterpret_cast<short *>(buf + o_mol);
        col.status = reinterpret_cast<signed char *>(buf + o_status);
        hitran_parse_kernel<<<(unsigned)((c + 127)/128), 128, 0, st>>>(
            buf, reinterpret_cast<uint64_t const *>(buf + o_start), c, col);
        e = cudaGetLastError();
        if (e != cudaSuccess) { rc = fail("cuda: %s (hitran_parse_kernel)", cudaGetErrorString(e)); break; }
        if (to_host(vnn, col.vnn, c, st) || to_host(snn, col.snn, c, st) ||
            to_host(yair, col.yair, c, st) || to_host(yself, col.yself, c, st) ||
            to_host(en, col.en, c, st) || to_host(n, col.n, c, st) || to_host(d, col.d, c, st) ||
            to_host(iso, col.iso, c, st) || to_host(mol, col.mol, c, st) ||
            to_host(status, col.status, c, st))
        {
            rc = GRT_SM100_ERR;
            break;
        }
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = fail("cuda: %s (hitran_parse_kernel)", cudaGetErrorString(e));
    } while (0);
    cudaFree(buf);
    return rc;
}

int grt_decimal_fields(int device, char const *text, int width, uint64_t count, double *out,
                       signed char *status)
{
    if (text == nullptr || out == nullptr || status == nullptr) return fail("null argument");
    if (width < 1 || width > 31) return fail("field width %d outside [1, 31]", width);
    if (count == 0) return 0;
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    GRT_CUDA(cudaSetDevice(device));
    char *buf = nullptr;
    double *o = nullptr;
    signed char *s = nullptr;
    GRT_CUDA(cudaMalloc(&buf, count*(uint64_t)width));
    GRT_CUDA(cudaMalloc(&o, 8*count));
    GRT_CUDA(cudaMalloc(&s, count));
    GRT_CUDA(cudaMemcpyAsync(buf, text, count*(uint64_t)width, cudaMemcpyHostToDevice, st));
    decimal_fields_kernel<<<(unsigned)((count + 127)/128), 128, 0, st>>>(buf, width, count, o, s);
    GRT_CUDA(cudaGetLastError());
    GRT_CUDA(cudaMemcpyAsync(out, o, 8*count, cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaMemcpyAsync(status, s, count, cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    GRT_CUDA(cudaFree(buf));
    GRT_CUDA(cudaFree(o));
    GRT_CUDA(cudaFree(s));
    return 0;
}

}  // extern "C"
