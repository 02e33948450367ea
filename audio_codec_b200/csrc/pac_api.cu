// Context creation: every read-only table the kernels use is built here on the
// host (long-double libm, rounded once to the working type) and uploaded once.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "pac_api.h"
#include "quant.cuh"

static const double kPi = 3.14159265358979323846;

template <typename T> static int upload(pac_ctx* c, const std::vector<T>& h, const T** out) {
    void* d = nullptr;
    if (cudaMalloc(&d, h.size() * sizeof(T) + 16) != cudaSuccess) return PAC_ERR_ALLOC;
    c->allocs.push_back(d);
    if (cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) return PAC_ERR_ALLOC;
    *out = (const T*)d;
    return 0;
}

// ---- windows (coder/window.py:14-92); arguments are formed in double exactly as
// NumPy forms them, then the transcendental is evaluated in long double.
static double w_sine(int n, int N) { return (double)sinl((long double)(kPi * (n + 0.5) / N)); }
static double w_hann(int n, int N) { return 0.5 * (1 - (double)cosl((long double)(2 * kPi * (n + 0.5) / N))); }
static long double bessel_i0(long double x) {
    long double s = 1, t = 1, q = x * x / 4;
    for (int k = 1; k < 200; ++k) {
        t *= q / ((long double)k * k);
        s += t;
        if (t < s * 1e-22L) break;
    }
    return s;
}
static double w_kaiser(int n, int N, double alpha) {
    double r = (2.0 * n + 1) / N - 1;
    double arg = kPi * alpha * sqrt(1 - r * r);
    return (double)(bessel_i0(arg) / bessel_i0(kPi * alpha));
}
static void make_windows(int N, int Ns, std::vector<double>& w) {
    w.assign((size_t)PAC_WIN_COUNT * N, 0.0);
    double* sine = &w[PAC_WIN_SINE * N];
    for (int n = 0; n < N; ++n) sine[n] = w_sine(n, N);
    for (int n = 0; n < N; ++n) w[PAC_WIN_KAISER * N + n] = w_kaiser(n, N, 4.0);
    if (N <= Ns) return;
    int pad = N / 4 - Ns / 4;
    double* st = &w[PAC_WIN_START * N];
    for (int n = 0; n < N / 2; ++n) st[n] = sine[n];
    for (int n = 0; n < pad; ++n) st[N / 2 + n] = 1.0;
    for (int n = 0; n < Ns / 2; ++n) st[N / 2 + pad + n] = w_sine(Ns / 2 + n, Ns);
    for (int n = 0; n < N; ++n) w[PAC_WIN_STOP * N + n] = st[N - 1 - n];
    double* ss = &w[PAC_WIN_STARTSTOP * N];
    for (int n = 0; n < Ns / 2; ++n) ss[pad + n] = w_sine(n, Ns);
    for (int n = 0; n < 2 * pad; ++n) ss[pad + Ns / 2 + n] = 1.0;
    for (int n = 0; n < Ns / 2; ++n) ss[3 * pad + Ns / 2 + n] = w_sine(Ns / 2 + n, Ns);
}

// ---- scale factor bands (coder/psychoac.py:100-160, coder/sbr.py:6-9)
static void make_bands(int nLines, int fs, PacBands* b) {
    static const double lim[25] = {100, 200, 300, 400, 510, 630, 770, 920, 1080, 1270, 1480, 1720, 2000, 2320,
                                   2700, 3150, 3700, 4400, 5300, 6400, 7700, 9500, 12000, 15500, 24000};
    double binw = (double)fs / (2 * nLines);
    double c[25], a[25];
    for (int i = 0; i < 25; ++i) c[i] = floor(lim[i] / binw - 0.5);
    for (int i = 0; i < 25; ++i) a[i] = c[i] - (i ? c[i - 1] : -1.0);
    for (int i = 0; i < 25; ++i)
        if (lim[i] > fs / 2.0) {
            double s = 0;
            for (int j = 0; j < i; ++j) s += a[j];
            a[i] = nLines - s;
            for (int j = i + 1; j < 25; ++j) a[j] = 0;
            break;
        }
    std::vector<int> nl;
    for (int i = 0; i < 25; ++i) nl.push_back((int)a[i]);
    size_t i = 1;
    while (i < nl.size()) {
        if (nl[i - 1] <= 12) {
            nl[i] += nl[i - 1];
            nl.erase(nl.begin() + (i - 1));
        } else {
            ++i;
        }
    }
    memset(b, 0, sizeof(*b));
    b->nBands = (int)nl.size();
    int lo = 0;
    for (int k = 0; k < b->nBands; ++k) {
        b->nLines[k] = nl[k];
        b->lower[k] = lo;
        lo += nl[k];
    }
}

// Window of any length for the window.* drop-ins (coder/window.py:14-92), cached per context.
int pac_window_table(pac_ctx* c, int kind, int N, int Ns, double alpha, const double** d, const float** f) {
    if (N < 1 || kind < 0 || kind > PAC_W_HANN || kind == PAC_W_NONE) return PAC_ERR_ARG;
    const bool edler = kind == PAC_W_START || kind == PAC_W_STOP || kind == PAC_W_STARTSTOP;
    if (edler && (Ns < 2 || Ns > N || (N % 4) || (Ns % 4))) return PAC_ERR_ARG;
    if (!edler) Ns = 0;
    if (kind != PAC_W_KAISER) alpha = 0.0;
    for (auto& e : c->winCache)
        if (e.kind == kind && e.N == N && e.Ns == Ns && e.alpha == alpha) { *d = e.d; *f = e.f; return 0; }
    std::vector<double> w(N, 0.0);
    const int pad = N / 4 - Ns / 4;
    switch (kind) {
        case PAC_W_SINE: for (int n = 0; n < N; ++n) w[n] = w_sine(n, N); break;
        case PAC_W_HANN: for (int n = 0; n < N; ++n) w[n] = w_hann(n, N); break;
        case PAC_W_KAISER: for (int n = 0; n < N; ++n) w[n] = w_kaiser(n, N, alpha); break;
        case PAC_W_START:
        case PAC_W_STOP:
            for (int n = 0; n < N / 2; ++n) w[n] = w_sine(n, N);
            for (int n = 0; n < pad; ++n) w[N / 2 + n] = 1.0;
            for (int n = 0; n < Ns / 2; ++n) w[N / 2 + pad + n] = w_sine(Ns / 2 + n, Ns);
            if (kind == PAC_W_STOP) for (int n = 0; n < N / 2; ++n) std::swap(w[n], w[N - 1 - n]);
            break;
        case PAC_W_STARTSTOP:
            for (int n = 0; n < Ns / 2; ++n) w[pad + n] = w_sine(n, Ns);
            for (int n = 0; n < 2 * pad; ++n) w[pad + Ns / 2 + n] = 1.0;
            for (int n = 0; n < Ns / 2; ++n) w[3 * pad + Ns / 2 + n] = w_sine(Ns / 2 + n, Ns);
            break;
    }
    std::vector<float> wf(w.begin(), w.end());
    pac_ctx::WinEntry e = {kind, N, Ns, alpha, nullptr, nullptr};
    int rc;
    if ((rc = upload(c, w, &e.d))) return rc;
    if ((rc = upload(c, wf, &e.f))) return rc;
    c->winCache.push_back(e);
    *d = e.d;
    *f = e.f;
    return 0;
}

template <typename T> static void to_type(const std::vector<long double>& s, std::vector<T>& d) {
    d.resize(s.size());
    for (size_t i = 0; i < s.size(); ++i) d[i] = (T)s[i];
}

template <typename T> static int make_numeric_tables(pac_ctx* c, PacTables<T>* t) {
    std::vector<double> wl, ws;
    make_windows(PAC_NLONG, PAC_NSHORT, wl);
    make_windows(PAC_NSHORT, PAC_NSHORT, ws);
    std::vector<T> tmp;
    int rc;
    auto up = [&](const std::vector<double>& v, const T** out) {
        tmp.assign(v.begin(), v.end());
        return upload(c, tmp, out);
    };
    if ((rc = up(wl, &t->win_long))) return rc;
    if ((rc = up(ws, &t->win_short))) return rc;
    std::vector<double> h(PAC_NLONG);
    for (int n = 0; n < PAC_NLONG; ++n) h[n] = w_hann(n, PAC_NLONG);
    if ((rc = up(h, &t->hann_long))) return rc;
    h.resize(PAC_NSHORT);
    for (int n = 0; n < PAC_NSHORT; ++n) h[n] = w_hann(n, PAC_NSHORT);
    if ((rc = up(h, &t->hann_short))) return rc;
    const long double pi = acosl(-1.0L);
    auto twid = [&](int N, const T** out) {   // exp(-j pi (8n+1)/(8M)), n < N/4
        int M = N / 2, Q = N / 4;
        std::vector<double> v(2 * Q);
        for (int n = 0; n < Q; ++n) {
            long double a = -pi * (8 * n + 1) / (8.0L * M);
            v[2 * n] = (double)cosl(a);
            v[2 * n + 1] = (double)sinl(a);
        }
        return up(v, out);
    };
    if ((rc = twid(PAC_NLONG, &t->tw_mdct_long))) return rc;
    if ((rc = twid(PAC_NSHORT, &t->tw_mdct_short))) return rc;
    // exp(-2 pi j k/n), k < count.  n divides 2048: every table is read off ONE first quadrant of the 2048-point roots by
    // symmetry (times (-j)^quadrant), so that the tables agree with each other entry for entry and a kernel that keeps only
    // a quadrant (fft.cuh: tw_quarter) reproduces them bit for bit; axis points are exact.
    std::vector<double> qc(513), qs(513);                  // cos, sin of 2 pi i / 2048, i <= 512
    for (int i = 0; i <= 512; ++i) {
        qc[i] = (double)cosl(2 * pi * i / 2048);
        qs[i] = (double)sinl(2 * pi * i / 2048);
    }
    qc[512] = 0.0; qs[512] = 1.0; qc[0] = 1.0; qs[0] = 0.0;
    auto roots = [&](int n, int count, const T** out) {
        std::vector<double> v(2 * count);
        for (int k = 0; k < count; ++k) {
            const int m = (int)(((long long)k * (2048 / n)) % 2048), qd = m / 512, i = m % 512;
            const double re = qc[i], im = -qs[i];          // quadrant 0
            double x, y;
            switch (qd) {
                case 0: x = re; y = im; break;
                case 1: x = im; y = -re; break;            // times -j
                case 2: x = -re; y = -im; break;
                default: x = -im; y = re; break;           // times +j
            }
            v[2 * k] = x;
            v[2 * k + 1] = y;
        }
        return up(v, out);
    };
    const int sizes[4] = {64, 128, 512, 1024};
    for (int i = 0; i < 4; ++i)
        if ((rc = roots(sizes[i], sizes[i], &t->tw_fft[i]))) return rc;
    if ((rc = roots(PAC_NLONG, PAC_NLONG / 2 + 1, &t->rtw_long))) return rc;
    if ((rc = roots(PAC_NSHORT, PAC_NSHORT / 2 + 1, &t->rtw_short))) return rc;
    return 0;
}

// ---- psychoacoustic per-line constants (coder/psychoac.py:35-48, :170-186)
static double bark_of(double f) {
    double a = (double)atanl((long double)(0.76 * f / 1000.0));
    double q = f / 7500.0;
    double b = (double)atanl((long double)(q * q));
    return 13.0 * a + 3.5 * b;
}
static double thresh_of(double f) {
    if (f < 10) f = 10;
    double u = f / 1000;
    double t1 = 3.64 * (double)powl((long double)u, -0.8L);
    double d = u - 3.3;
    double t2 = 6.5 * (double)expl((long double)(-0.6 * (d * d)));
    double t3 = 0.001 * (double)powl((long double)u, 4.0L);
    return t1 - t2 + t3;
}
static int make_psy(pac_ctx* c, int half, const double** bark, const double** thr, const double** ff,
                    double* norm) {
    int fs = c->cfg.sample_rate, N = 2 * half, rc;
    std::vector<double> b(half), t(half), f(half + 1);
    double spacing = (double)fs / (2 * half);
    for (int k = 0; k < half; ++k) {
        double fr = spacing * (k + 0.5);
        b[k] = bark_of(fr);
        t[k] = thresh_of(fr);
    }
    double val = 1.0 / (N * (1.0 / fs));
    for (int k = 0; k <= half; ++k) f[k] = k * val;
    long double acc = 0;
    const long double pi = acosl(-1.0L);
    for (int n = 0; n < N; ++n) {
        long double hn = 0.5L - 0.5L * cosl(2 * pi * n / (N - 1));
        acc += hn * hn;
    }
    *norm = (double)(4.0L / ((long double)N * N * (acc / N)));
    if ((rc = upload(c, b, bark))) return rc;
    if ((rc = upload(c, t, thr))) return rc;
    return upload(c, f, ff);
}

// ---- PVQ codebook sizes N(L,K) (coder/gain_shape_quantize.py:69-101):
// N(L,K) = N(L-1,K) + N(L-1,K-1) + N(L,K-1), N(L,0) = 1, N(0,K>=1) = 0.
// Row L >= 4 is stored up to and including its first entry above 2^32; rows
// 0..3 have closed forms (0/1, 2, 4K, 4K^2+2) and are not stored.
static int make_pvq(pac_ctx* c) {
    const unsigned long long LIM = 1ull << 32;
    std::vector<unsigned long long> all;
    std::vector<int> off(PAC_PVQ_LMAX + 2, 0);
    std::vector<unsigned long long> prev, cur;
    // row 3 as a function
    auto row3 = [](unsigned long long k) { return k == 0 ? 1ull : 4ull * k * k + 2ull; };
    for (int L = 4; L <= PAC_PVQ_LMAX; ++L) {
        cur.clear();
        cur.push_back(1);
        for (size_t k = 1;; ++k) {
            unsigned long long up0 = (L == 4) ? row3(k) : (k < prev.size() ? prev[k] : ~0ull);
            unsigned long long up1 = (L == 4) ? row3(k - 1) : prev[k - 1];
            if (up0 == ~0ull) { cur.push_back(LIM + 1); break; }   // parent already over the cap
            unsigned long long v = up0 + up1 + cur[k - 1];
            cur.push_back(v);
            if (v > LIM) break;
        }
        off[L] = (int)all.size();
        all.insert(all.end(), cur.begin(), cur.end());
        prev = cur;
    }
    off[PAC_PVQ_LMAX + 1] = (int)all.size();
    for (int L = 0; L < 4; ++L) off[L] = 0;
    int rc;
    std::vector<double> lg(PAC_PVQ_LMAX + 1, 0.0);
    for (int L = 1; L <= PAC_PVQ_LMAX; ++L) lg[L] = (double)log2l((long double)L);
    {   // K(L, bits): largest K with N(L,K) <= 2^bits, and ceil(log2 N(L,K)) with the 1-bit minimum
        // (coder/gain_shape_quantize.py:275-291)
        std::vector<int2> kw((size_t)(PAC_PVQ_LMAX + 1) * 33);
        auto NLK = [&](int L, unsigned long long k) -> unsigned long long {
            if (k == 0) return 1ull;
            if (L <= 0) return 0ull;
            if (L == 1) return 2ull;
            if (L == 2) return 4ull * k;
            if (L == 3) return 4ull * k * k + 2ull;
            size_t o = off[L], len = off[L + 1] - off[L];
            return k >= len ? (1ull << 40) : all[o + k];
        };
        for (int L = 0; L <= PAC_PVQ_LMAX; ++L)
            for (int b = 0; b <= 32; ++b) {
                unsigned long long lim = 1ull << b, k = 0;
                if (L == 2) k = lim / 4ull;
                else if (L == 3) { while (NLK(3, k + 1) <= lim) ++k; }
                else if (L >= 4) {
                    size_t o = off[L], len = off[L + 1] - off[L];
                    size_t lo = 0, hi = len;
                    while (lo < hi) { size_t mid = (lo + hi) / 2; if (all[o + mid] > lim) hi = mid; else lo = mid + 1; }
                    k = lo - 1;
                }
                unsigned long long n = NLK(L, k);
                int w = 1;
                if (n > 1) { w = 0; unsigned long long v = n - 1; while (v) { ++w; v >>= 1; } }
                kw[(size_t)L * 33 + b] = make_int2((int)k, w);
            }
        if ((rc = upload(c, kw, &c->host.pvqKW))) return rc;
    }
    if ((rc = upload(c, lg, &c->host.log2tab))) return rc;
    if ((rc = upload(c, all, &c->host.pvqN))) return rc;
    return upload(c, off, &c->host.pvqRowOff);
}

// ---- quantised-angle table (PacDev::thetaTab)
__global__ void k_theta_tab(double* tab) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < 2 || e >= (2 << PAC_THETA_BITS)) return;
    const int aT = 31 - __clz(e);
    const unsigned long long code = (unsigned long long)(e - (1 << aT));
    const double theta = dequantize_uniform(code, aT) * 1.5707963267948966;      // * np.pi / 2
    double sn, ct;
    pm_sincos(theta, &sn, &ct);
    tab[4 * e + 0] = theta;
    tab[4 * e + 1] = pm_log2(pm_tan(fabs(theta)) + PAC_EPS);
    tab[4 * e + 2] = ct;
    tab[4 * e + 3] = sn;
}
static int make_theta(pac_ctx* c) {
    void* d = nullptr;
    const size_t n = (size_t)(2 << PAC_THETA_BITS) * 4;
    if (cudaMalloc(&d, n * sizeof(double)) != cudaSuccess) return PAC_ERR_ALLOC;
    c->allocs.push_back(d);
    if (cudaMemset(d, 0, n * sizeof(double)) != cudaSuccess) return PAC_ERR_ALLOC;
    k_theta_tab<<<(2 << PAC_THETA_BITS) / 256, 256>>>((double*)d);
    if (cudaDeviceSynchronize() != cudaSuccess) return PAC_ERR_ALLOC;
    c->host.thetaTab = (const double*)d;
    return 0;
}

// ---- SBR decode: scipy.ndimage.gaussian_filter1d(sigma=200, mode='reflect',
// truncate=4) response to a unit impulse at each of the first nOm positions of a
// length-2*nLen signal (coder/codec.py:139-146); only the first nLen outputs used.
static int make_sbr(pac_ctx* c) {
    PacBands& b = c->host.longBands;
    int cut = -1, nOm = 0;
    int upper_last = b.lower[b.nBands - 1] + b.nLines[b.nBands - 1] - 1;
    for (int k = 0; k < b.nBands; ++k) {
        b.omitted[k] = (c->cfg.use_sbr && b.lower[k] >= upper_last / 2) ? 1 : 0;
        if (b.lower[k] >= upper_last / 2) {
            if (cut < 0) cut = b.lower[k];
            ++nOm;
        }
    }
    c->host.sbrCutoff = cut;
    c->host.sbrNOmitted = nOm;
    int nLen = PAC_HALF_LONG - cut, n2 = 2 * nLen;
    const double sigma = 200.0;
    int lw = (int)(4.0 * sigma + 0.5);
    std::vector<double> ker(2 * lw + 1);
    double s = 0;
    for (int i = -lw; i <= lw; ++i) s += (ker[i + lw] = exp(-0.5 / (sigma * sigma) * (double)i * i));
    for (auto& v : ker) v /= s;
    std::vector<double> g((size_t)nOm * nLen, 0.0);
    for (int p = 0; p < nOm; ++p)
        for (int i = 0; i < nLen; ++i) {
            double acc = 0;
            for (int j = -lw; j <= lw; ++j) {
                // reflect extension (d c b a | a b c d | d c b a) of index i+j
                long long q = i + j;
                long long per = 2LL * n2;
                q = ((q % per) + per) % per;
                if (q >= n2) q = per - 1 - q;
                if (q == p) acc += ker[j + lw];
            }
            g[(size_t)p * nLen + i] = acc;
        }
    return upload(c, g, &c->host.sbrGauss);
}

extern "C" int pac_create(int device, const pac_config* cfg, pac_ctx** out) {
    if (!cfg || !out) return PAC_ERR_ARG;
    if (cfg->n_mdct_lines != PAC_HALF_LONG || cfg->n_channels < 1 || cfg->sample_rate < 8000) return PAC_ERR_ARG;
    if (cfg->n_scale_bits < 1 || cfg->n_scale_bits > 8 || cfg->n_mant_size_bits < 1 || cfg->n_mant_size_bits > 16)
        return PAC_ERR_ARG;
    if (cfg->precision != PAC_F32 && cfg->precision != PAC_F64) return PAC_ERR_ARG;
    {   // the device quantisers hold R = 2^nScaleBits - 1 + nMantBits magnitude bits in 64-bit integers
        // (quant.cuh); the reference uses Python bigints.  Same bound as pac_scale_factor.
        const int maxMant = (1 << cfg->n_mant_size_bits) > 16 ? 16 : (1 << cfg->n_mant_size_bits);
        if ((1 << cfg->n_scale_bits) - 1 + (maxMant > 5 ? maxMant : 5) > 62) return PAC_ERR_UNSUPPORTED;
    }
    if (cudaSetDevice(device) != cudaSuccess) return PAC_ERR_ARG;
    pac_ctx* c = new pac_ctx();
    c->device = device;
    c->optBatchCb = 0;
    c->optSmemPad = 0;
    c->optHostChunkMb = 0.0;
    c->optVariant = 0;
    c->cfg = *cfg;
    c->dev = nullptr;
    c->scratch = nullptr;
    c->scratchBytes = 0;
    c->pinned = nullptr;
    c->pinnedBytes = 0;
    for (int i = 0; i < 4; ++i) cudaEventCreate(&c->ev[i]);
    for (int k = 0; k < 4; ++k) {
        c->kevN[k] = 0;
        for (int j = 0; j < PAC_KEV_MAX; ++j) { cudaEventCreate(&c->kev[k][j][0]); cudaEventCreate(&c->kev[k][j][1]); }
    }
    c->launches = 0;
    c->evValid[0] = c->evValid[1] = 0;
    c->io[0] = c->io[1] = nullptr;
    c->ioBytes[0] = c->ioBytes[1] = 0;
    memset(&c->host, 0, sizeof(c->host));
    PacDev& h = c->host;
    h.fs = cfg->sample_rate;
    h.nChannels = cfg->n_channels;
    h.nScaleBits = cfg->n_scale_bits;
    h.nMantSizeBits = cfg->n_mant_size_bits;
    h.useSBR = cfg->use_sbr;
    h.useVQ = cfg->use_vq;
    h.maxMantBits = (1 << cfg->n_mant_size_bits) > 16 ? 16 : (1 << cfg->n_mant_size_bits);
    h.targetBitsPerSample = cfg->target_bits_per_sample;
    make_bands(PAC_HALF_LONG, cfg->sample_rate, &h.longBands);
    make_bands(PAC_HALF_SHORT, cfg->sample_rate, &h.shortBands);
    int rc = 0;
    if (h.longBands.nBands > PAC_MAX_BANDS) rc = PAC_ERR_ARG;
    if (!rc) rc = make_numeric_tables<double>(c, &h.d);
    if (!rc) rc = make_numeric_tables<float>(c, &h.f);
    if (!rc) rc = make_psy(c, PAC_HALF_LONG, &h.barkLong, &h.threshLong, &h.fftFreqLong, &h.fftNormLong);
    if (!rc) rc = make_psy(c, PAC_HALF_SHORT, &h.barkShort, &h.threshShort, &h.fftFreqShort, &h.fftNormShort);
    h.fftValLong = 1.0 / (PAC_NLONG * (1.0 / cfg->sample_rate));     // as make_psy
    h.fftValShort = 1.0 / (PAC_NSHORT * (1.0 / cfg->sample_rate));
    for (int w = 0; w < 2 && !rc; ++w) {          // band of every MDCT line
        const PacBands& bd = w ? h.shortBands : h.longBands;
        std::vector<unsigned char> lb(w ? PAC_HALF_SHORT : PAC_HALF_LONG, 0);
        for (int b = 0; b < bd.nBands; ++b)
            for (int i = 0; i < bd.nLines[b]; ++i)
                if (bd.lower[b] + i < (int)lb.size()) lb[bd.lower[b] + i] = (unsigned char)b;
        rc = upload(c, lb, w ? &h.lineBandShort : &h.lineBandLong);
    }
    if (!rc) rc = make_pvq(c);
    if (!rc) rc = make_sbr(c);
    if (!rc) rc = make_theta(c);
    if (!rc) rc = pac_ops_static_setup(device);     // this device is current (cudaSetDevice above)
    if (!rc) {
        void* d = nullptr;
        if (cudaMalloc(&d, sizeof(PacDev)) != cudaSuccess) rc = PAC_ERR_ALLOC;
        else {
            c->allocs.push_back(d);
            c->dev = (PacDev*)d;
            if (cudaMemcpy(d, &h, sizeof(PacDev), cudaMemcpyHostToDevice) != cudaSuccess) rc = PAC_ERR_ALLOC;
        }
    }
    if (rc) {
        pac_destroy(c);
        return rc;
    }
    *out = c;
    return 0;
}

extern "C" void pac_destroy(pac_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (void* p : c->allocs) cudaFree(p);
    if (c->scratch) cudaFree(c->scratch);
    if (c->pinned) cudaFreeHost(c->pinned);
    for (int i = 0; i < 2; ++i) if (c->io[i]) cudaFree(c->io[i]);
    for (int i = 0; i < 4; ++i) cudaEventDestroy(c->ev[i]);
    for (int k = 0; k < 4; ++k)
        for (int j = 0; j < PAC_KEV_MAX; ++j) { cudaEventDestroy(c->kev[k][j][0]); cudaEventDestroy(c->kev[k][j][1]); }
    if (c->cpReady)
    {
        for (int i = 0; i < 2; ++i) { cudaStreamDestroy(c->cp[i]); cudaEventDestroy(c->cpEv[i]); }
        for (int i = 0; i < 8; ++i) cudaEventDestroy(c->chunkEv[i]);
    }
    delete c;
}

int pac_scratch(pac_ctx* c, size_t bytes, void** out) {
    if (bytes > c->scratchBytes) {
        if (c->scratch) cudaFree(c->scratch);
        c->scratch = nullptr;
        c->scratchBytes = 0;
        if (cudaMalloc(&c->scratch, bytes) != cudaSuccess) return PAC_ERR_ALLOC;
        c->scratchBytes = bytes;
    }
    *out = c->scratch;
    return 0;
}

int pac_io(pac_ctx* c, int which, size_t bytes, void** out) {
    if (bytes > c->ioBytes[which]) {
        if (c->io[which]) cudaFree(c->io[which]);
        c->io[which] = nullptr;
        c->ioBytes[which] = 0;
        if (cudaMalloc(&c->io[which], bytes + 256) != cudaSuccess) return PAC_ERR_ALLOC;
        c->ioBytes[which] = bytes;
    }
    *out = c->io[which];
    return 0;
}

int pac_pinned(pac_ctx* c, size_t bytes, void** out) {
    if (bytes > c->pinnedBytes) {
        if (c->pinned) cudaFreeHost(c->pinned);
        c->pinned = nullptr;
        c->pinnedBytes = 0;
        if (cudaMallocHost(&c->pinned, bytes) != cudaSuccess) return PAC_ERR_ALLOC;
        c->pinnedBytes = bytes;
    }
    *out = c->pinned;
    return 0;
}

extern "C" const char* pac_strerror(int s) {
    switch (s) {
        case PAC_OK: return "ok";
        case PAC_ERR_ARG: return "bad argument";
        case PAC_ERR_ALLOC: return "allocation failed";
        case PAC_ERR_UNSUPPORTED: return "unsupported configuration (undefined in the reference)";
        case PAC_ERR_SOFT: return "per-frame soft error flagged on device";
        case PAC_ERR_TOO_LARGE: return "request exceeds the fixed static scratch of a no-allocation entry point";
        default:
            if (s <= -1000) return cudaGetErrorString((cudaError_t)(-s - 1000));
            return "unknown";
    }
}

extern "C" int pac_get_bands(const pac_ctx* c, int which, int* n_lines, int* omitted) {
    if (!c) return PAC_ERR_ARG;
    const PacBands& b = which ? c->host.shortBands : c->host.longBands;
    for (int k = 0; k < b.nBands; ++k) {
        if (n_lines) n_lines[k] = b.nLines[k];
        if (omitted) omitted[k] = b.omitted[k];
    }
    return b.nBands;
}

// Tuning knobs and test hooks (batch size of the thread-per-band kernels, shared-memory padding,
// host pipeline chunk size).  None of them changes a result.
extern "C" int pac_set_option(pac_ctx* c, int option, double value) {
    if (!c || value < 0.0) return PAC_ERR_ARG;
    switch (option) {
        case PAC_OPT_BATCH_CB: c->optBatchCb = (long long)value; return 0;
        case PAC_OPT_SMEM_PAD: c->optSmemPad = (int)value; return 0;
        case PAC_OPT_HOST_CHUNK_MB: c->optHostChunkMb = value; return 0;
        case PAC_OPT_VARIANT: c->optVariant = (int)value; return 0;
        default: return PAC_ERR_ARG;
    }
}

// Device time of the dominant kernel of the last pac_encode_streams / pac_decode_streams
// call on this context (CUDA events recorded on the caller's stream around the launch).
extern "C" int pac_last_kernel_ms(pac_ctx* c, float* enc_ms, float* dec_ms) {
    if (!c) return PAC_ERR_ARG;
    if (enc_ms) { *enc_ms = -1.f; if (c->evValid[0]) cudaEventElapsedTime(enc_ms, c->ev[0], c->ev[1]); }
    if (dec_ms) { *dec_ms = -1.f; if (c->evValid[1]) cudaEventElapsedTime(dec_ms, c->ev[2], c->ev[3]); }
    return 0;
}

// Device time and launch count of one kernel over the last pac_encode_streams / pac_decode_streams call
// (kind: 0 k_encode, 1 k_pvq, 2 k_dequant, 3 k_decode); the call must have completed.
extern "C" int pac_kernel_stats(pac_ctx* c, int kind, float* ms, int* launches) {
    if (!c || kind < 0 || kind > 3) return PAC_ERR_ARG;
    float tot = 0.f;
    const int n = c->kevN[kind] < PAC_KEV_MAX ? c->kevN[kind] : PAC_KEV_MAX;
    for (int j = 0; j < n; ++j) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, c->kev[kind][j][0], c->kev[kind][j][1]) != cudaSuccess) return PAC_ERR_ARG;
        tot += t;
    }
    if (ms) *ms = tot;
    if (launches) *launches = c->kevN[kind];
    return 0;
}

// Kernels launched by the library on this context since pac_create.
extern "C" long long pac_kernel_launches(pac_ctx* c) { return c ? c->launches : PAC_ERR_ARG; }
