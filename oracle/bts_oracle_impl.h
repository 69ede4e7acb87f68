/*
 * bts_oracle_impl.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference hot path (reduction_1x1 + decoder glue +
 * local_planar_guidance of /root/reference/pytorch/bts.py) including the analytic backward that
 * PyTorch autograd derives for it. Included twice by bts_oracle.c with REAL = float / double and
 * SUF = f32 / f64. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may call it. The product (bts_b200/) never links or imports it.
 *
 * Parity pin: tests/test_oracle_golden.py checks every function below against the vectors in
 * tests/golden/ (npz files), which were produced by running the unmodified reference in the build
 * container (tests/golden/make_golden.py).
 *
 * Reference lines followed (all in pytorch/bts.py):
 *   chain construction / order of layers ............ 98-122
 *   ELU after every inter_* conv, none after last ... 112-119
 *   sigmoid head (is_final) ......................... 108-110
 *   theta/phi/dist scaling and n1..n4 ............... 127-134
 *   F.normalize(n[:, :3], 2, 1) + cat ............... 250-253
 *   LPG u/v grids, den, abs_min, clamp, division .... 150-173
 */

#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

/* ELU with alpha = 1 as ATen computes it: x > 0 ? x : exp(x) - 1   (bts.py:116-119) */
static REAL FN(elu)(REAL t) { return t > (REAL)0 ? t : (REAL)(EXP(t) - (REAL)1); }
static REAL FN(sigm)(REAL t) { return (REAL)1 / ((REAL)1 + EXP(-t)); }

/*
 * Chain forward for one pixel. act holds every layer's post-activation vector back to back
 * (act[0..dims[0]) = input, ...). Returns pointer to the last layer's raw logits (3 or 1 values).
 * W[l] is PyTorch's [out,in,1,1] weight viewed as row-major [dims[l+1]][dims[l]].
 */
static void FN(chain_pixel_fwd)(const REAL *const *W, const int *dims, int n_layers, REAL *act)
{
    REAL *a = act;
    for (int l = 0; l < n_layers; ++l) {
        const int cin = dims[l], cout = dims[l + 1];
        REAL *z = a + cin;
        for (int o = 0; o < cout; ++o) {
            REAL s = 0;
            for (int c = 0; c < cin; ++c) s += W[l][(size_t)o * cin + c] * a[c];
            z[o] = (l + 1 < n_layers) ? FN(elu)(s) : s; /* last conv has no ELU (bts.py:112-113) */
        }
        a = z;
    }
}

/*
 * reduction_1x1.forward (bts.py:124-136).
 *   x   [B][dims[0]][hw]   NCHW, hw = h*w
 *   out [B][4][hw]  (n1,n2,n3,n4)   or [B][1][hw] sigmoid(logit) when is_final
 */
void FN(oracle_reduc_fwd)(const REAL *x, const REAL *const *W, const int *dims, int n_layers, int is_final,
                          REAL max_depth, int B, int hw, REAL *out)
{
    int tot = 0;
    for (int l = 0; l <= n_layers; ++l) tot += dims[l];
    REAL *act = (REAL *)malloc(sizeof(REAL) * (size_t)tot);
    const REAL pi = (REAL)M_PI;
    for (int b = 0; b < B; ++b)
        for (int p = 0; p < hw; ++p) {
            for (int c = 0; c < dims[0]; ++c) act[c] = x[((size_t)b * dims[0] + c) * hw + p];
            FN(chain_pixel_fwd)(W, dims, n_layers, act);
            const REAL *z = act + tot - dims[n_layers];
            if (is_final) {
                out[(size_t)b * hw + p] = FN(sigm)(z[0]);
            } else {
                const REAL theta = FN(sigm)(z[0]) * pi / (REAL)3;        /* bts.py:127 */
                const REAL phi = FN(sigm)(z[1]) * pi * (REAL)2;          /* bts.py:128 */
                const REAL dist = FN(sigm)(z[2]) * max_depth;            /* bts.py:129 */
                REAL *o = out + (size_t)b * 4 * hw + p;
                o[0] = SIN(theta) * COS(phi);                            /* bts.py:130 */
                o[(size_t)hw] = SIN(theta) * SIN(phi);                   /* bts.py:131 */
                o[(size_t)2 * hw] = COS(theta);                          /* bts.py:132 */
                o[(size_t)3 * hw] = dist;                                /* bts.py:133 */
            }
        }
    free(act);
}

/*
 * Backward of reduction_1x1.forward. grad_out has the shape of out. grad_x like x; grad_W[l] like
 * W[l], ACCUMULATED into (caller zeroes). Autograd-equivalent formulas, SURVEY.md 8(a) row a8.
 */
void FN(oracle_reduc_bwd)(const REAL *x, const REAL *const *W, const int *dims, int n_layers, int is_final,
                          REAL max_depth, int B, int hw, const REAL *grad_out, REAL *grad_x,
                          REAL *const *grad_W)
{
    int tot = 0, cmax = 0;
    for (int l = 0; l <= n_layers; ++l) {
        tot += dims[l];
        if (dims[l] > cmax) cmax = dims[l];
    }
    REAL *act = (REAL *)malloc(sizeof(REAL) * (size_t)tot);
    REAL *d0 = (REAL *)malloc(sizeof(REAL) * (size_t)cmax);
    REAL *d1 = (REAL *)malloc(sizeof(REAL) * (size_t)cmax);
    const REAL pi = (REAL)M_PI;
    for (int b = 0; b < B; ++b)
        for (int p = 0; p < hw; ++p) {
            for (int c = 0; c < dims[0]; ++c) act[c] = x[((size_t)b * dims[0] + c) * hw + p];
            FN(chain_pixel_fwd)(W, dims, n_layers, act);
            const REAL *z = act + tot - dims[n_layers];
            REAL *delta = d0, *dprev = d1;
            if (is_final) {
                const REAL s = FN(sigm)(z[0]);
                delta[0] = grad_out[(size_t)b * hw + p] * s * ((REAL)1 - s);
            } else {
                const REAL s0 = FN(sigm)(z[0]), s1 = FN(sigm)(z[1]), s2 = FN(sigm)(z[2]);
                const REAL theta = s0 * pi / (REAL)3, phi = s1 * pi * (REAL)2;
                const REAL st = SIN(theta), ct = COS(theta), sp = SIN(phi), cp = COS(phi);
                const REAL *g = grad_out + (size_t)b * 4 * hw + p;
                const REAL g1 = g[0], g2 = g[(size_t)hw], g3 = g[(size_t)2 * hw], g4 = g[(size_t)3 * hw];
                const REAL dtheta = g1 * ct * cp + g2 * ct * sp - g3 * st;
                const REAL dphi = -g1 * st * sp + g2 * st * cp;
                delta[0] = dtheta * (pi / (REAL)3) * s0 * ((REAL)1 - s0);
                delta[1] = dphi * (pi * (REAL)2) * s1 * ((REAL)1 - s1);
                delta[2] = g4 * max_depth * s2 * ((REAL)1 - s2);
            }
            /* walk the chain backwards: a_in of layer l starts at offset sum(dims[0..l)) */
            int off_out = tot - dims[n_layers];
            for (int l = n_layers - 1; l >= 0; --l) {
                const int cin = dims[l], cout = dims[l + 1];
                const int off_in = off_out - cin;
                const REAL *a_in = act + off_in;
                for (int o = 0; o < cout; ++o)
                    for (int c = 0; c < cin; ++c) grad_W[l][(size_t)o * cin + c] += delta[o] * a_in[c];
                for (int c = 0; c < cin; ++c) {
                    REAL s = 0;
                    for (int o = 0; o < cout; ++o) s += W[l][(size_t)o * cin + c] * delta[o];
                    if (l > 0) {
                        /* ELU'(t) = t > 0 ? 1 : exp(t) = a + 1 where a = ELU(t) <= 0 */
                        const REAL a = a_in[c];
                        s *= (a > (REAL)0) ? (REAL)1 : (a + (REAL)1);
                    }
                    dprev[c] = s;
                }
                REAL *t = delta; delta = dprev; dprev = t;
                off_out = off_in;
            }
            for (int c = 0; c < dims[0]; ++c) grad_x[((size_t)b * dims[0] + c) * hw + p] = delta[c];
        }
    free(act); free(d0); free(d1);
}

/* decoder glue bts.py:250-253: n_hat = n / max(||n||_2, 1e-12); plane_eq = cat(n_hat, n4) */
void FN(oracle_glue_fwd)(const REAL *out4, int B, int hw, REAL *plane_eq)
{
    for (int b = 0; b < B; ++b)
        for (int p = 0; p < hw; ++p) {
            const REAL *s = out4 + (size_t)b * 4 * hw + p;
            REAL *d = plane_eq + (size_t)b * 4 * hw + p;
            const REAL n1 = s[0], n2 = s[(size_t)hw], n3 = s[(size_t)2 * hw];
            REAL nrm = SQRT(n1 * n1 + n2 * n2 + n3 * n3);
            if (nrm < (REAL)1e-12) nrm = (REAL)1e-12;
            d[0] = n1 / nrm; d[(size_t)hw] = n2 / nrm; d[(size_t)2 * hw] = n3 / nrm;
            d[(size_t)3 * hw] = s[(size_t)3 * hw];
        }
}

/* backward of the glue: G_n = (G - n_hat (n_hat . G)) / ||n||   (clamp_min inactive for unit n) */
void FN(oracle_glue_bwd)(const REAL *out4, const REAL *grad_plane_eq, int B, int hw, REAL *grad_out4)
{
    for (int b = 0; b < B; ++b)
        for (int p = 0; p < hw; ++p) {
            const REAL *s = out4 + (size_t)b * 4 * hw + p;
            const REAL *g = grad_plane_eq + (size_t)b * 4 * hw + p;
            REAL *d = grad_out4 + (size_t)b * 4 * hw + p;
            const REAL n1 = s[0], n2 = s[(size_t)hw], n3 = s[(size_t)2 * hw];
            REAL nrm = SQRT(n1 * n1 + n2 * n2 + n3 * n3);
            if (nrm < (REAL)1e-12) nrm = (REAL)1e-12;
            const REAL h1 = n1 / nrm, h2 = n2 / nrm, h3 = n3 / nrm;
            const REAL g1 = g[0], g2 = g[(size_t)hw], g3 = g[(size_t)2 * hw];
            const REAL dot = h1 * g1 + h2 * g2 + h3 * g3;
            d[0] = (g1 - h1 * dot) / nrm; d[(size_t)hw] = (g2 - h2 * dot) / nrm;
            d[(size_t)2 * hw] = (g3 - h3 * dot) / nrm; d[(size_t)3 * hw] = g[(size_t)3 * hw];
        }
}

/*
 * local_planar_guidance.forward (bts.py:149-173). plane_eq [B][4][h][w] -> depth [B][h*r][w*r].
 * Operation order follows the reference exactly -- (n1*u + n2*v) + n3 with separately rounded
 * products, then n4 / den -- so the f32 instantiation is bit-identical to the reference on CPU.
 * abs_min = min |den| BEFORE the clamp (bts.py:167). focal is not an input: the reference ignores it.
 */
void FN(oracle_lpg_fwd)(const REAL *plane_eq, int B, int h, int w, int r, REAL *depth, REAL *abs_min)
{
    const REAL eps = (REAL)1e-3;                                           /* bts.py:168 */
    const size_t hw = (size_t)h * w;
    const int H = h * r, Wd = w * r;
    REAL amin = (REAL)INFINITY;
    int have_nan = 0;
    for (int b = 0; b < B; ++b)
        for (int y = 0; y < H; ++y)
            for (int xx = 0; xx < Wd; ++xx) {
                const int i = y / r, j = xx / r;                           /* repeat_interleave, bts.py:150-151 */
                const REAL *pe = plane_eq + (size_t)b * 4 * hw + (size_t)i * w + j;
                const REAL u = ((REAL)(xx % r) - (REAL)(r - 1) * (REAL)0.5) / (REAL)r; /* bts.py:157-158 */
                const REAL v = ((REAL)(y % r) - (REAL)(r - 1) * (REAL)0.5) / (REAL)r;  /* bts.py:160-161 */
                volatile REAL t1 = pe[0] * u;            /* volatile: forbid FMA contraction */
                volatile REAL t2 = pe[hw] * v;
                volatile REAL t3 = t1 + t2;
                REAL den = t3 + pe[2 * hw];                                /* bts.py:166 */
                const REAL ad = FABS(den);
                if (ad != ad) have_nan = 1;
                if (ad < amin) amin = ad;                                  /* bts.py:167 */
                if (den > (REAL)0 && den < eps) den = eps;                 /* bts.py:170 */
                if (den < (REAL)0 && den > -eps) den = -eps;               /* bts.py:171 */
                depth[((size_t)b * H + y) * Wd + xx] = pe[3 * hw] / den;   /* bts.py:173 */
            }
    *abs_min = have_nan ? (REAL)NAN : amin;                                /* torch.min propagates NaN */
}

/*
 * Backward of LPG: grad_plane_eq [B][4][h][w] from grad_depth [B][H][W].
 *   unclamped pixel: dG_k = -n4 * g * w_k / den^2, w = (u, v, 1);   clamped pixel: 0 for k=1..3
 *   dG_4 = g / den_c always.   (torch.where routes no gradient into the replaced branch.)
 */
void FN(oracle_lpg_bwd)(const REAL *grad_depth, const REAL *plane_eq, int B, int h, int w, int r,
                        REAL *grad_plane_eq)
{
    const REAL eps = (REAL)1e-3;
    const size_t hw = (size_t)h * w;
    const int H = h * r, Wd = w * r;
    for (int b = 0; b < B; ++b)
        for (int i = 0; i < h; ++i)
            for (int j = 0; j < w; ++j) {
                const REAL *pe = plane_eq + (size_t)b * 4 * hw + (size_t)i * w + j;
                const REAL n1 = pe[0], n2 = pe[hw], n3 = pe[2 * hw], n4 = pe[3 * hw];
                REAL G1 = 0, G2 = 0, G3 = 0, G4 = 0;
                for (int p = 0; p < r; ++p)
                    for (int q = 0; q < r; ++q) {
                        const REAL u = ((REAL)q - (REAL)(r - 1) * (REAL)0.5) / (REAL)r;
                        const REAL v = ((REAL)p - (REAL)(r - 1) * (REAL)0.5) / (REAL)r;
                        const REAL g = grad_depth[((size_t)b * H + (size_t)i * r + p) * Wd + (size_t)j * r + q];
                        REAL den = n1 * u + n2 * v + n3;
                        int clamped = 0;
                        if (den > (REAL)0 && den < eps) { den = eps; clamped = 1; }
                        if (den < (REAL)0 && den > -eps) { den = -eps; clamped = 1; }
                        G4 += g / den;
                        if (!clamped) {
                            const REAL t = -n4 * g / (den * den);
                            G1 += t * u; G2 += t * v; G3 += t;
                        }
                    }
                REAL *d = grad_plane_eq + (size_t)b * 4 * hw + (size_t)i * w + j;
                d[0] = G1; d[hw] = G2; d[2 * hw] = G3; d[3 * hw] = G4;
            }
}

#undef FN
#undef CAT
#undef CAT_
