"""An INDEPENDENT numpy/scipy statement of what cellector computes (written from the algorithm's definition:
beta-binomial log pmf via scipy.special.gammaln/betaln, numpy sorting for the R-8 quantiles), used to pin the
C++ oracle: it shares no code and no numerics (Lanczos vs scipy's lgamma) with oracle/.  Small inputs only."""
from __future__ import annotations

import numpy as np
from scipy.special import betaln, gammaln


def log_betabinom(a, r, alpha, beta):
    n = a + r
    lbc = gammaln(n + 1) - gammaln(a + 1) - gammaln(r + 1)
    return lbc + betaln(a + alpha, r + beta) - betaln(alpha, beta)


def quantile_r8(x, tau):
    x = np.sort(np.asarray(x, float))
    n = len(x)
    h = (n + 1.0 / 3.0) * tau + 1.0 / 3.0
    hf = int(h)
    if hf <= 0:
        return x[0]
    if hf >= n:
        return x[-1]
    return x[hf - 1] + (h - hf) * (x[min(hf, n - 1)] - x[hf - 1])


def median(x):
    x = np.sort(np.asarray(x, float))
    n = len(x)
    return x[n // 2] if n % 2 else 0.5 * (x[n // 2 - 1] + x[n // 2])


class PyRef:
    def __init__(self, locus, cell, ref, alt, n_loci0, n_cells, min_ref=4, min_alt=4):
        locus, cell, ref, alt = (np.asarray(v, np.int64) for v in (locus, cell, ref, alt))
        cnt_ref = np.bincount(locus[ref > 0], minlength=n_loci0)
        cnt_alt = np.bincount(locus[alt > 0], minlength=n_loci0)
        used0 = (cnt_ref >= min_ref) & (cnt_alt >= min_alt)
        self.locus_ids = np.flatnonzero(used0)
        remap = -np.ones(n_loci0, np.int64)
        remap[self.locus_ids] = np.arange(len(self.locus_ids))
        keep = used0[locus]
        self.l, self.c, self.r, self.a = remap[locus[keep]], cell[keep], ref[keep], alt[keep]
        self.L, self.N = len(self.locus_ids), n_cells
        self.tot_ref = np.bincount(self.l, self.r, self.L)
        self.tot_alt = np.bincount(self.l, self.a, self.L)
        self.used = np.ones(self.L, bool)
        self.excluded = np.zeros(n_cells, bool)
        self.row_len = np.bincount(self.c, minlength=n_cells)

    def alpha_beta(self, excl):
        m = excl[self.c]
        alpha = self.tot_alt + 1 - np.bincount(self.l[m], self.a[m], self.L)
        beta = self.tot_ref + 1 - np.bincount(self.l[m], self.r[m], self.L)
        return alpha, beta

    def score(self, alpha, beta, used):
        v = log_betabinom(self.a, self.r, alpha[self.l], beta[self.l])
        m = used[self.l]
        ll = np.bincount(self.c[m], v[m], self.N)
        n = np.bincount(self.c[m], minlength=self.N)
        return ll, n, v

    def iteration(self, iqr_multiple=5.0):
        alpha, beta = self.alpha_beta(self.excluded)
        ll, n, v = self.score(alpha, beta, self.used)
        norm = np.where(n > 0, ll / np.maximum(n, 1), 0.0)
        q1, q3 = quantile_r8(norm, 0.25), quantile_r8(norm, 0.75)
        thr = q1 - iqr_multiple * (q3 - q1)
        new = norm < thr
        m = self.used[self.l]
        emin = m & new[self.c]
        emaj = m & ~new[self.c]
        out = dict(ll=ll, n_loci=n, norm=norm, median=median(norm), q1=q1, q3=q3, threshold=thr, excluded=new,
                   n_new=int((new & ~self.excluded).sum()), n_rescued=int((~new & self.excluded).sum()),
                   ll_minority=np.bincount(self.l[emin], v[emin], self.L), ll_majority=np.bincount(self.l[emaj], v[emaj], self.L),
                   cells_minority=np.bincount(self.l[emin], minlength=self.L), cells_majority=np.bincount(self.l[emaj], minlength=self.L),
                   ref_minority=np.bincount(self.l[emin], self.r[emin], self.L), alt_minority=np.bincount(self.l[emin], self.a[emin], self.L),
                   ref_majority=np.bincount(self.l[emaj], self.r[emaj], self.L), alt_majority=np.bincount(self.l[emaj], self.a[emaj], self.L))
        cm = out["cells_minority"]
        per_cell = np.where(cm > 0, out["ll_minority"] / np.maximum(cm, 1), 0.0)
        self.used = self.used & ~(per_cell < -80.0)
        out["loci_used"] = self.used.copy()
        self.excluded = new
        return out

    def posteriors(self, posterior_threshold=0.999, min_loci_used=30):
        ex = self.excluded
        a_maj, b_maj = self.alpha_beta(ex)
        a_min, b_min = self.alpha_beta(~ex)
        mf = (ex.sum() + 1.0) / (self.N + 1.0)
        a_dbl = (a_maj - 1) * mf + (a_min - 1) + 1
        b_dbl = (b_maj - 1) * mf + (b_min - 1) + 1
        mf = max(mf, 0.01)
        a_maj, b_maj = (a_maj - 1) * mf + 1, (b_maj - 1) * mf + 1
        allu = np.ones(self.L, bool)
        ll_min = self.score(a_min, b_min, allu)[0]
        ll_maj = self.score(a_maj, b_maj, allu)[0]
        ll_dbl = self.score(a_dbl, b_dbl, allu)[0]
        lp_d = np.log(self.N / 1000.0 / 100.0 * max(mf, 0.1))
        lnum = np.log(mf) + ll_min
        lden = np.logaddexp(np.logaddexp(lnum, np.log(1 - mf) + ll_maj), lp_d + ll_dbl)
        post, dpost = np.exp(lnum - lden), np.exp(lp_d + ll_dbl - lden)
        label = np.full(self.N, 3)
        label[post > posterior_threshold] = 0
        label[(post <= posterior_threshold) & (1 - post > posterior_threshold)] = 1
        label[dpost > 0.5] = 2
        label[self.row_len < min_loci_used] = 3
        return dict(posterior=post, doublet_posterior=dpost, ll_major=ll_maj, ll_minor=ll_min, label=label,
                    anomaly=np.where(ex, 0, 1))
