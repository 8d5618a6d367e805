// niftyb200 -- the O(n_bins) amplitude (power-spectrum) chain of the correlated field and its VJP, on the device.
//
// Reference: non_parametric_amplitude.correlate (src/re/correlated_field.py:277-304) with the integrated Wiener
// process (src/re/gauss_markov.py:105-121) and _remove_slope (:144-147); numpy twin _Amplitude / _Normalization
// / _TwoLogIntegrations / _SlopeRemover (src/library/correlated_fields.py:91-201,276-385).  Runs once per
// nonlinear evaluation (not per CG iteration); on the device so that the n_bins-sized tables never travel to
// the host and back.  All arithmetic in fp64.
//
// With n = nb - 2 steps, a_m = flx sqrt(dt_m), q_m = dt_m^2 / 12 + asp:
//   dev_m = a_m xi[m][1]                       lvl_m = a_m xi[m][0] sqrt(q_m) + dt_m dev_m / 2
//   rate_m = sum_{k<m} dev_k                   inc_m = lvl_m + dt_m rate_m
//   twolog_0 = twolog_1 = 0, twolog_{m+2} = sum_{k<=m} inc_k
//   ln_k = slope rel_k + twolog_k - twolog_{nb-1} rel_k / rel_{nb-1}          spec_k = exp(ln_k)
//   power kind:      S = sum_{k>=1} mult_k spec_k,    amp_k = c sqrt(spec_k)
//   amplitude kind:  S = sum_{k>=1} mult_k spec_k^2,  amp_k = c spec_k
//   c = flu * totvol / (sqrt(S) * azm)   (= flu sqrt(totvol) / norm / azm, norm = sqrt(S / totvol));  amp_0 = totvol
// (amp_k, k >= 1, is the NORMALISED amplitude amp / azm of get_normalized_amplitudes, :605-611.)
//
// One CTA; thread t owns the contiguous segment [t seg, (t+1) seg) of the steps, so every prefix sum is
// "segment sums -> scan of the T partials by thread 0 -> rescan of the segment": three barrier-separated phases.
#pragma once

#include "nb2_common.cuh"

namespace nb2 {

struct AmpGeom {
    long long nb;          // number of power bins (>= 1)
    const double* rel;     // relative_log_mode_lengths (nb)
    const double* dt;      // log_volume (nb - 2)
    const double* mult;    // mode_multiplicity as fp64 (nb)
    double totvol;
};
struct AmpScalars {
    double flu, slope, flx, asp, azm;   // prior-transformed values; flx < 0: no flexibility; asp < 0: no asperity
    int power_kind;                     // 1: "power" parametrisation, 0: "amplitude"
};

constexpr int kAmpThreads = 512;

NB2_HD void amp_segment(long long n, int tid, int nthreads, long long& s0, long long& s1) {
    const long long seg = (n + nthreads - 1) / nthreads;
    s0 = seg * tid < n ? seg * tid : n;
    s1 = s0 + seg < n ? s0 + seg : n;
}

struct AmplitudeForward {
    static constexpr int NPHASES = 7;
    static constexpr int MAX_THREADS = kAmpThreads;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        AmpGeom g;
        AmpScalars s;
        const double* xi;      // (nb - 2, 2) white excitations of the Wiener process (may be NULL without flexibility)
        double* inc;           // scratch (nb - 2)
        double* ln;            // scratch (nb): ln_k, kept for the VJP
        double* sums;          // out: sums[0] = S, sums[1] = c
        double* amp;           // out (nb): normalised amplitude, amp[0] = totvol
        void* table;           // optional: table[table_idx[k]] = amp[k] in `table_f32 ? float : double` (ksq table)
        const long long* table_idx;
        int table_f32;
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads + 2) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long, int tid, int nthreads, char* smem_raw) {
        double* sm = reinterpret_cast<double*>(smem_raw);      // T partials, then [T] = total, [T+1] = c
        const long long nb = p.g.nb, n = nb > 2 ? nb - 2 : 0;
        const bool flex = p.s.flx >= 0.0 && n > 0 && p.xi != nullptr;
        const double asp = p.s.asp >= 0.0 ? p.s.asp : 0.0;
        long long s0, s1;
        amp_segment(n, tid, nthreads, s0, s1);
        if (PH == 0) {          // segment sums of dev
            double acc = 0.0;
            if (flex)
                for (long long m = s0; m < s1; ++m) acc += p.s.flx * sqrt(p.g.dt[m]) * p.xi[2 * m + 1];
            sm[tid] = acc;
        } else if (PH == 1 || PH == 3) {   // thread 0: exclusive scan of the partials, total behind them
            if (tid == 0) {
                double run = 0.0;
                for (int t = 0; t < nthreads; ++t) {
                    const double v = sm[t];
                    sm[t] = run;
                    run += v;
                }
                sm[nthreads] = run;
            }
        } else if (PH == 2) {   // rate (prefix of dev) -> inc, segment sums of inc
            double rate = sm[tid], acc = 0.0;
            if (flex)
                for (long long m = s0; m < s1; ++m) {
                    const double dt = p.g.dt[m], a = p.s.flx * sqrt(dt);
                    const double dev = a * p.xi[2 * m + 1];
                    const double lvl = a * p.xi[2 * m] * sqrt(dt * dt / 12.0 + asp) + 0.5 * dt * dev;
                    const double inc = lvl + dt * rate;
                    p.inc[m] = inc;
                    acc += inc;
                    rate += dev;
                }
            sm[tid] = acc;
        } else if (PH == 4) {   // twolog (prefix of inc) -> ln_k, segment sums of S
            const double total = sm[nthreads];           // twolog_{nb-1}
            const double rlast = nb > 1 ? p.g.rel[nb - 1] : 1.0;
            double tw = sm[tid], acc = 0.0;
            auto bin = [&](long long k, double twolog) {
                const double r = p.g.rel[k];
                const double ln = p.s.slope * r + (flex ? twolog - total * (r / rlast) : 0.0);
                p.ln[k] = ln;
                const double sp = exp(ln);
                acc += p.g.mult[k] * (p.s.power_kind ? sp : sp * sp);
            };
            if (tid == 0) {
                p.ln[0] = 0.0;
                if (nb > 1) bin(1, 0.0);
            }
            for (long long m = s0; m < s1; ++m) {
                if (flex) tw += p.inc[m];
                bin(m + 2, tw);
            }
            sm[tid] = acc;
        } else if (PH == 5) {
            if (tid == 0) {
                double S = 0.0;
                for (int t = 0; t < nthreads; ++t) S += sm[t];
                const double c = p.s.flu * p.g.totvol / (sqrt(S) * p.s.azm);
                sm[nthreads] = S;
                sm[nthreads + 1] = c;
                p.sums[0] = S;
                p.sums[1] = c;
            }
        } else if (PH == 6) {
            const double c = sm[nthreads + 1];
            auto put = [&](long long k, double v) {
                p.amp[k] = v;
                if (p.table) {
                    if (p.table_f32) reinterpret_cast<float*>(p.table)[p.table_idx[k]] = float(v);
                    else reinterpret_cast<double*>(p.table)[p.table_idx[k]] = v;
                }
            };
            if (tid == 0) {
                put(0, p.g.totvol);
                if (nb > 1) put(1, c * exp((p.s.power_kind ? 0.5 : 1.0) * p.ln[1]));
            }
            for (long long m = s0; m < s1; ++m) put(m + 2, c * exp((p.s.power_kind ? 0.5 : 1.0) * p.ln[m + 2]));
        }
    }
};

// ---------------------------------------------------------------- VJP of the chain
// Given ampbar (cotangent of the normalised amplitudes; entry 0 is a constant and ignored) and the forward's
// ln / amp / S, returns xibar (nb - 2, 2) and scalars_bar = (flubar, slopebar, flxbar, aspbar, azmbar).
//   G = sum_{k>=1} ampbar_k amp_k;   flubar = G / flu;   azmbar = -G / azm
//   power:      lb_k = ampbar_k amp_k / 2 - G mult_k spec_k / (2 S)
//   amplitude:  lb_k = ampbar_k amp_k     - G mult_k spec_k^2 / S                 (cotangent of ln_k, k >= 1)
//   slopebar = sum lb_k rel_k;  twologbar_k = lb_k - [k = nb-1] slopebar / rel_{nb-1}
//   incbar_m = sum_{m' >= m} twologbar_{m'+2}                                     (reverse inclusive scan)
//   devbar_m = sum_{m' > m} dt_{m'} incbar_{m'} + dt_m incbar_m / 2               (reverse exclusive scan)
//   xibar[m][1] = a_m devbar_m;  xibar[m][0] = a_m sqrt(q_m) incbar_m
//   flxbar = sum (xi[m][1] devbar_m + xi[m][0] sqrt(q_m) incbar_m) sqrt(dt_m)
//   aspbar = sum a_m xi[m][0] incbar_m / (2 sqrt(q_m))
struct AmplitudeVjp {
    static constexpr int NPHASES = 10;
    static constexpr int MAX_THREADS = kAmpThreads;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        AmpGeom g;
        AmpScalars s;
        const double* xi;
        const double* amp;      // forward output (nb)
        const double* sums;     // forward output: S, c
        const double* ampbar;   // (nb)
        const double* ln;       // forward scratch (nb): ln_k, read only
        double* lb;             // work (nb): cotangent of ln_k
        double* inc;            // work (nb - 2): incbar_m
        double* xibar;          // out (nb - 2, 2) (may be NULL without flexibility)
        double* scalars_bar;    // out (5)
    };
    static size_t smem_bytes(int nthreads) { return size_t(2 * nthreads + 4) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }

    // thread 0: suffix (reverse exclusive) scan of the T partials
    NB2_HD static void suffix_scan(double* sm, int nthreads) {
        double run = 0.0;
        for (int t = nthreads - 1; t >= 0; --t) {
            const double v = sm[t];
            sm[t] = run;
            run += v;
        }
    }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long, int tid, int nthreads, char* smem_raw) {
        double* sm = reinterpret_cast<double*>(smem_raw);
        double* sm2 = sm + nthreads;
        double* sc = sm + 2 * nthreads;      // [0] = G, [1] = slopebar
        const long long nb = p.g.nb, n = nb > 2 ? nb - 2 : 0;
        const bool flex = p.s.flx >= 0.0 && n > 0 && p.xi != nullptr;
        const double asp = p.s.asp >= 0.0 ? p.s.asp : 0.0;
        long long s0, s1;
        amp_segment(n, tid, nthreads, s0, s1);
        if (PH == 0) {          // G
            double acc = 0.0;
            if (tid == 0 && nb > 1) acc += p.ampbar[1] * p.amp[1];
            for (long long m = s0; m < s1; ++m) acc += p.ampbar[m + 2] * p.amp[m + 2];
            sm[tid] = acc;
        } else if (PH == 1) {
            if (tid == 0) {
                double G = 0.0;
                for (int t = 0; t < nthreads; ++t) G += sm[t];
                sc[0] = G;
                p.scalars_bar[0] = G / p.s.flu;
                p.scalars_bar[4] = -G / p.s.azm;
            }
        } else if (PH == 2) {   // lb_k, slopebar partials
            const double G = sc[0], S = p.sums[0];
            double acc = 0.0;
            auto bin = [&](long long k) {
                const double sp = exp(p.ln[k]);
                const double aa = p.ampbar[k] * p.amp[k];
                const double lb = p.s.power_kind ? 0.5 * aa - 0.5 * G * p.g.mult[k] * sp / S
                                                 : aa - G * p.g.mult[k] * sp * sp / S;
                p.lb[k] = lb;
                acc += lb * p.g.rel[k];
            };
            if (tid == 0 && nb > 1) bin(1);
            for (long long m = s0; m < s1; ++m) bin(m + 2);
            sm[tid] = acc;
        } else if (PH == 3) {
            if (tid == 0) {
                double sb = 0.0;
                for (int t = 0; t < nthreads; ++t) sb += sm[t];
                sc[1] = sb;
                p.scalars_bar[1] = sb;
                p.scalars_bar[2] = 0.0;
                p.scalars_bar[3] = 0.0;
            }
        } else if (PH == 4) {   // segment sums of twologbar_{m+2}
            double acc = 0.0;
            if (flex) {
                const double corr = -sc[1] / p.g.rel[nb - 1];
                for (long long m = s0; m < s1; ++m) acc += p.lb[m + 2] + (m + 2 == nb - 1 ? corr : 0.0);
            }
            sm[tid] = acc;
        } else if (PH == 5 || PH == 7) {
            if (tid == 0) suffix_scan(PH == 5 ? sm : sm2, nthreads);
        } else if (PH == 6) {   // incbar (reverse inclusive scan), segment sums of dt incbar
            double suf = sm[tid], acc = 0.0;
            if (flex) {
                const double corr = -sc[1] / p.g.rel[nb - 1];
                for (long long m = s1 - 1; m >= s0; --m) {
                    suf += p.lb[m + 2] + (m + 2 == nb - 1 ? corr : 0.0);
                    p.inc[m] = suf;
                    acc += p.g.dt[m] * suf;
                }
            }
            sm2[tid] = acc;
        } else if (PH == 8) {   // devbar (reverse exclusive scan of dt incbar), xibar, flx / asp partials
            double suf = sm2[tid], fa = 0.0, aa = 0.0;
            if (flex) {
                for (long long m = s1 - 1; m >= s0; --m) {
                    const double dt = p.g.dt[m], ib = p.inc[m];
                    const double devbar = suf + 0.5 * dt * ib;
                    suf += dt * ib;
                    const double sq = sqrt(dt), a = p.s.flx * sq, rq = sqrt(dt * dt / 12.0 + asp);
                    const double x0 = p.xi[2 * m], x1 = p.xi[2 * m + 1];
                    p.xibar[2 * m] = a * rq * ib;
                    p.xibar[2 * m + 1] = a * devbar;
                    fa += (x1 * devbar + x0 * rq * ib) * sq;
                    aa += a * x0 * ib / (2.0 * rq);
                }
            }
            sm[tid] = fa;
            sm2[tid] = aa;
        } else if (PH == 9) {
            if (tid == 0) {
                double fa = 0.0, aa = 0.0;
                for (int t = 0; t < nthreads; ++t) {
                    fa += sm[t];
                    aa += sm2[t];
                }
                p.scalars_bar[2] = fa;
                p.scalars_bar[3] = p.s.asp >= 0.0 ? aa : 0.0;
            }
        }
    }
};

// ---------------------------------------------------------------- JVP of the chain (forward mode)
// Tangents (dxi (nb - 2, 2), dflu, dslope, dflx, dasp, dazm) -> damp (nb), the tangent of the normalised
// amplitudes (damp_0 = 0), with the forward's ln / amp / S:
//   da_m = dflx sqrt(dt_m);  ddev_m = da_m xi1 + a_m dxi1
//   dlvl_m = (da_m xi0 + a_m dxi0) sqrt(q_m) + a_m xi0 dasp / (2 sqrt(q_m)) + dt_m ddev_m / 2
//   drate / dinc / dtwolog: the same two prefix sums as the forward;  dln_k = dslope rel_k + dtwolog_k - dtwolog_last rel_k / rel_last
//   power:      dS = sum mult spec dln,      damp_k = amp_k (dc/c + dln_k / 2)
//   amplitude:  dS = sum 2 mult spec^2 dln,  damp_k = amp_k (dc/c + dln_k)
//   dc/c = dflu / flu - dazm / azm - dS / (2 S)
struct AmplitudeJvp {
    static constexpr int NPHASES = 7;
    static constexpr int MAX_THREADS = kAmpThreads;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        AmpGeom g;
        AmpScalars s;
        AmpScalars ds;          // tangents of (flu, slope, flx, asp, azm); power_kind unused
        const double* xi;
        const double* dxi;      // (nb - 2, 2) (NULL: zero)
        const double* ln;       // forward scratch (nb), read only
        const double* amp;      // forward output
        const double* sums;     // forward: S, c
        double* dinc;           // scratch (nb - 2)
        double* dln;            // scratch (nb)
        double* damp;           // out (nb)
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads + 2) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }

    NB2_HD static void step(const Params& p, long long m, double asp, double& ddev, double& dlvl) {
        const double dt = p.g.dt[m], sq = sqrt(dt), a = p.s.flx * sq, da = p.ds.flx * sq;
        const double x0 = p.xi[2 * m], x1 = p.xi[2 * m + 1];
        const double d0 = p.dxi ? p.dxi[2 * m] : 0.0, d1 = p.dxi ? p.dxi[2 * m + 1] : 0.0;
        const double rq = sqrt(dt * dt / 12.0 + asp);
        ddev = da * x1 + a * d1;
        dlvl = (da * x0 + a * d0) * rq + (p.s.asp >= 0.0 ? a * x0 * p.ds.asp / (2.0 * rq) : 0.0) + 0.5 * dt * ddev;
    }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long, int tid, int nthreads, char* smem_raw) {
        double* sm = reinterpret_cast<double*>(smem_raw);
        const long long nb = p.g.nb, n = nb > 2 ? nb - 2 : 0;
        const bool flex = p.s.flx >= 0.0 && n > 0 && p.xi != nullptr;
        const double asp = p.s.asp >= 0.0 ? p.s.asp : 0.0;
        long long s0, s1;
        amp_segment(n, tid, nthreads, s0, s1);
        if (PH == 0) {
            double acc = 0.0;
            if (flex)
                for (long long m = s0; m < s1; ++m) {
                    double ddev, dlvl;
                    step(p, m, asp, ddev, dlvl);
                    acc += ddev;
                }
            sm[tid] = acc;
        } else if (PH == 1 || PH == 3) {
            if (tid == 0) {
                double run = 0.0;
                for (int t = 0; t < nthreads; ++t) {
                    const double v = sm[t];
                    sm[t] = run;
                    run += v;
                }
                sm[nthreads] = run;
            }
        } else if (PH == 2) {
            double rate = sm[tid], acc = 0.0;
            if (flex)
                for (long long m = s0; m < s1; ++m) {
                    double ddev, dlvl;
                    step(p, m, asp, ddev, dlvl);
                    const double dinc = dlvl + p.g.dt[m] * rate;
                    p.dinc[m] = dinc;
                    acc += dinc;
                    rate += ddev;
                }
            sm[tid] = acc;
        } else if (PH == 4) {
            const double total = sm[nthreads];
            const double rlast = nb > 1 ? p.g.rel[nb - 1] : 1.0;
            double tw = sm[tid], acc = 0.0;
            auto bin = [&](long long k, double dtw) {
                const double r = p.g.rel[k];
                const double dln = p.ds.slope * r + (flex ? dtw - total * (r / rlast) : 0.0);
                p.dln[k] = dln;
                const double sp = exp(p.ln[k]);
                acc += p.g.mult[k] * (p.s.power_kind ? sp : 2.0 * sp * sp) * dln;
            };
            if (tid == 0) {
                p.dln[0] = 0.0;
                if (nb > 1) bin(1, 0.0);
            }
            for (long long m = s0; m < s1; ++m) {
                if (flex) tw += p.dinc[m];
                bin(m + 2, tw);
            }
            sm[tid] = acc;
        } else if (PH == 5) {
            if (tid == 0) {
                double dS = 0.0;
                for (int t = 0; t < nthreads; ++t) dS += sm[t];
                sm[nthreads] = p.ds.flu / p.s.flu - p.ds.azm / p.s.azm - 0.5 * dS / p.sums[0];   // dc / c
            }
        } else if (PH == 6) {
            const double dcc = sm[nthreads], f = p.s.power_kind ? 0.5 : 1.0;
            if (tid == 0) {
                p.damp[0] = 0.0;
                if (nb > 1) p.damp[1] = p.amp[1] * (dcc + f * p.dln[1]);
            }
            for (long long m = s0; m < s1; ++m) p.damp[m + 2] = p.amp[m + 2] * (dcc + f * p.dln[m + 2]);
        }
    }
};

}  // namespace nb2
