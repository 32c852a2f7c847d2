/* oracle/gl_oracle.c -- TEST INFRASTRUCTURE ONLY.  See gl_oracle.h.
 *
 * Plain-C restatement of the reference's commit path (Goldilocks NTT / LDE,
 * Poseidon sponge, binary Poseidon Merkle tree, Merkle proofs).  Written for
 * clarity, not speed: radix-2 transforms, one permutation at a time, OpenMP
 * over rows.  Used only by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg as the checker.
 */
#include "gl_oracle.h"
#include "poseidon_constants.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ field */

uint64_t glo_canon(uint64_t a) { return a >= GLO_P ? a - GLO_P : a; }

uint64_t glo_add(uint64_t a, uint64_t b)
{
    u128 s = (u128)glo_canon(a) + glo_canon(b);
    if (s >= GLO_P) s -= GLO_P;
    return (uint64_t)s;
}

uint64_t glo_sub(uint64_t a, uint64_t b)
{
    a = glo_canon(a);
    b = glo_canon(b);
    return a >= b ? a - b : a + (GLO_P - b);
}

uint64_t glo_mul(uint64_t a, uint64_t b) { return (uint64_t)(((u128)a * b) % GLO_P); }

uint64_t glo_pow(uint64_t a, uint64_t e)
{
    uint64_t r = 1;
    a = glo_canon(a);
    while (e) {
        if (e & 1) r = glo_mul(r, a);
        a = glo_mul(a, a);
        e >>= 1;
    }
    return r;
}

uint64_t glo_inv(uint64_t a) { return glo_pow(a, GLO_P - 2); }

/* Goldilocks::W[32] (goldilocks_base_field.cpp:44); W[k-1] = W[k]^2, W[0] = 1. */
uint64_t glo_root_of_unity(unsigned k)
{
    uint64_t w = 7277203076849721926ULL;
    for (unsigned i = 32; i > k; i--) w = glo_mul(w, w);
    return w;
}

static unsigned ilog2(uint64_t n)
{
    unsigned k = 0;
    while (((uint64_t)1 << k) < n) k++;
    return k;
}

int glo_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void glo_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

/* -------------------------------------------------------------------- NTT */

static uint64_t bitrev(uint64_t x, unsigned bits)
{
    uint64_t r = 0;
    for (unsigned i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}

/* In-place forward DFT of every column of the size x ncols row-major matrix a,
 * with primitive root w = W[log2 size]: Y[i] = sum_j X[j] w^(ij), natural order. */
static void dft_columns(uint64_t *a, uint64_t size, uint64_t ncols)
{
    unsigned k = ilog2(size);
    if (size <= 1) return;
    /* bit-reversal of rows (ntt_goldilocks.cpp:243-305 does the same permutation) */
    uint64_t *tmp = (uint64_t *)malloc(ncols * sizeof(uint64_t));
    for (uint64_t i = 0; i < size; i++) {
        uint64_t r = bitrev(i, k);
        if (r > i) {
            memcpy(tmp, a + i * ncols, ncols * 8);
            memcpy(a + i * ncols, a + r * ncols, ncols * 8);
            memcpy(a + r * ncols, tmp, ncols * 8);
        }
    }
    free(tmp);
    uint64_t *tw = (uint64_t *)malloc((size / 2) * sizeof(uint64_t));
    uint64_t w = glo_root_of_unity(k);
    tw[0] = 1;
    for (uint64_t i = 1; i < size / 2; i++) tw[i] = glo_mul(tw[i - 1], w);
    for (unsigned s = 1; s <= k; s++) {
        uint64_t m = (uint64_t)1 << s, half = m >> 1, step = size / m;
#pragma omp parallel for schedule(static)
        for (uint64_t blk = 0; blk < size / m; blk++) {
            for (uint64_t j = 0; j < half; j++) {
                uint64_t wj = tw[j * step];
                uint64_t *lo = a + (blk * m + j) * ncols;
                uint64_t *hi = lo + half * ncols;
                for (uint64_t c = 0; c < ncols; c++) {
                    uint64_t t = glo_mul(wj, hi[c]);
                    uint64_t u = lo[c];
                    lo[c] = glo_add(u, t);
                    hi[c] = glo_sub(u, t);
                }
            }
        }
    }
    free(tw);
}

void glo_ntt(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend)
{
    if (size == 0 || ncols == 0) return; /* ntt_goldilocks.cpp:168-171 */
    uint64_t *d = dst ? dst : (uint64_t *)src;
    if (d != src) memmove(d, src, size * ncols * 8);
    for (uint64_t i = 0; i < size * ncols; i++) d[i] = glo_canon(d[i]);
    dft_columns(d, size, ncols);
    if (!inverse) return;
    /* INTT = forward transform, then row x -> row (size - x) mod size, scaled by 1/size
     * (ntt_goldilocks.cpp:139-148) or by r_[row] = 7^row / size when extend (:124-136). */
    unsigned k = ilog2(size);
    uint64_t ninv = glo_inv((uint64_t)1 << k);
    if (size == 1) ninv = 1;
    uint64_t *rowtmp = (uint64_t *)malloc(ncols * 8);
    for (uint64_t x = 1; x < size - x; x++) { /* swap rows x and size-x */
        uint64_t y = size - x;
        memcpy(rowtmp, d + x * ncols, ncols * 8);
        memcpy(d + x * ncols, d + y * ncols, ncols * 8);
        memcpy(d + y * ncols, rowtmp, ncols * 8);
    }
    free(rowtmp);
    uint64_t sc = ninv;
    for (uint64_t r = 0; r < size; r++) {
        for (uint64_t c = 0; c < ncols; c++) d[r * ncols + c] = glo_mul(d[r * ncols + c], sc);
        if (extend) sc = glo_mul(sc, 7);
    }
}

void glo_extend_pol(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (n == 0 || ncols == 0) return;
    glo_ntt(out, in, n, ncols, 1, 1);                       /* ntt_goldilocks.cpp:345 */
    memset(out + n * ncols, 0, (n_ext - n) * ncols * 8);    /* zero-pad: :259-278 */
    glo_ntt(out, out, n_ext, ncols, 0, 0);                  /* :346 */
}

/* --------------------------------------------------------------- Poseidon */

static uint64_t pow7(uint64_t x)
{
    uint64_t x2 = glo_mul(x, x), x3 = glo_mul(x, x2), x4 = glo_mul(x2, x2);
    return glo_mul(x3, x4);
}

/* new[i] = sum_j mat(j,i) old[j]  (poseidon_goldilocks.hpp:148-162) */
static void mds_layer(uint64_t st[12])
{
    uint64_t old[12];
    memcpy(old, st, sizeof old);
    for (int i = 0; i < 12; i++) {
        u128 acc = 0; /* 12 * 64 * 2^64 < 2^128 */
        for (int j = 0; j < 12; j++) {
            uint64_t m = GLP_MDS_CIRC[(i - j + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0);
            acc += (u128)m * old[j];
        }
        st[i] = (uint64_t)(acc % GLO_P);
    }
}

static void dense_layer(uint64_t st[12])
{
    uint64_t old[12];
    memcpy(old, st, sizeof old);
    for (int i = 0; i < 12; i++) {
        uint64_t acc = 0;
        for (int j = 0; j < 12; j++) acc = glo_add(acc, glo_mul(GLP_PM[j * 12 + i], old[j]));
        st[i] = acc;
    }
}

void glo_hash_full_result(uint64_t state[12], const uint64_t input[12])
{
    uint64_t st[12];
    for (int i = 0; i < 12; i++) st[i] = glo_add(input[i], GLP_RC[i]);
    for (int r = 0; r < 3; r++) {
        for (int i = 0; i < 12; i++) st[i] = glo_add(pow7(st[i]), GLP_RC[(r + 1) * 12 + i]);
        mds_layer(st);
    }
    for (int i = 0; i < 12; i++) st[i] = glo_add(pow7(st[i]), GLP_RC[48 + i]);
    dense_layer(st);
    for (int r = 0; r < 22; r++) {
        const uint64_t *S = &GLP_SP[23 * r];
        uint64_t t = glo_add(pow7(st[0]), GLP_RC[60 + r]);
        uint64_t s0 = glo_mul(t, S[0]);
        for (int i = 1; i < 12; i++) s0 = glo_add(s0, glo_mul(st[i], S[i]));
        for (int i = 1; i < 12; i++) st[i] = glo_add(st[i], glo_mul(t, S[11 + i]));
        st[0] = s0;
    }
    for (int r = 0; r < 3; r++) {
        for (int i = 0; i < 12; i++) st[i] = glo_add(pow7(st[i]), GLP_RC[82 + 12 * r + i]);
        mds_layer(st);
    }
    for (int i = 0; i < 12; i++) st[i] = pow7(st[i]);
    mds_layer(st);
    memcpy(state, st, sizeof st);
}

void glo_linear_hash(uint64_t out[4], const uint64_t *in, uint64_t size)
{
    if (size <= 4) { /* returned unhashed: poseidon_goldilocks.cpp:43-48 */
        for (uint64_t i = 0; i < 4; i++) out[i] = i < size ? glo_canon(in[i]) : 0;
        return;
    }
    uint64_t st[12];
    uint64_t remaining = size;
    while (remaining) {
        if (remaining == size) memset(st + 8, 0, 4 * 8);
        else memcpy(st + 8, st, 4 * 8);
        uint64_t n = remaining < 8 ? remaining : 8;
        memset(st, 0, 8 * 8);
        memcpy(st, in + (size - remaining), n * 8);
        glo_hash_full_result(st, st);
        remaining -= n;
    }
    memcpy(out, st, 4 * 8);
}

uint64_t glo_tree_num_elements(uint64_t height) { return height * 4 + (height - 1) * 4; }

static void tree_levels(uint64_t *tree, uint64_t nrows)
{
    uint64_t pending = nrows, next_index = 0;
    while (pending > 1) {
        uint64_t next_n = (pending - 1) / 2 + 1;
#pragma omp parallel for schedule(static)
        for (uint64_t i = 0; i < next_n; i++) {
            uint64_t in[12], st[12];
            memset(in, 0, sizeof in);
            memcpy(in, &tree[next_index + i * 8], 8 * 8);
            glo_hash_full_result(st, in);
            memcpy(&tree[next_index + (pending + i) * 4], st, 4 * 8);
        }
        next_index += pending * 4;
        pending /= 2;
    }
}

void glo_merkletree(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows)
{
    if (nrows == 0) return;
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < nrows; i++) glo_linear_hash(&tree[i * 4], &input[i * ncols], ncols);
    tree_levels(tree, nrows);
}

void glo_merkletree_batch(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch)
{
    if (nrows == 0) return;
    uint64_t nb = ncols > 0 ? (ncols + batch - 1) / batch : 1;
    uint64_t nlast = ncols - (nb - 1) * batch;
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < nrows; i++) {
        uint64_t *buf = (uint64_t *)malloc(nb * 4 * 8);
        for (uint64_t j = 0; j < nb; j++)
            glo_linear_hash(&buf[j * 4], &input[i * ncols + j * batch], j == nb - 1 ? nlast : batch);
        glo_linear_hash(&tree[i * 4], buf, nb * 4);
        free(buf);
    }
    tree_levels(tree, nrows);
}

uint64_t glo_merkle_proof_size(uint64_t height)
{
    if (height > 1) return (uint64_t)ceil(log10((double)height) / log10(2.0)); /* merkleTreeGL.hpp:71-78 */
    return 0;
}

void glo_group_proof(uint64_t *proof, const uint64_t *source, const uint64_t *nodes,
                     uint64_t height, uint64_t width, uint64_t idx)
{
    for (uint64_t i = 0; i < width; i++) proof[i] = glo_canon(source[idx * width + i]);
    uint64_t *p = proof + width;
    uint64_t offset = 0, n = height * 4;
    while (n > 4) { /* merkleTreeGL.cpp:24-35, recursion unrolled */
        uint64_t si = (idx ^ 1) * 4;
        for (int k = 0; k < 4; k++) p[k] = glo_canon(nodes[offset + si + k]);
        p += 4;
        idx >>= 1;
        uint64_t next_n = ((n - 1) / 8 + 1) * 4;
        offset += next_n * 2;
        n = next_n;
    }
}

void glo_fill_xorshift(uint64_t *dst, uint64_t count, uint64_t *state)
{
    uint64_t s = *state;
    for (uint64_t i = 0; i < count; i++) {
        s ^= s << 13;
        s ^= s >> 7;
        s ^= s << 17;
        dst[i] = s % GLO_P;
    }
    *state = s;
}

/* ------------------------------------------------- next rows: Q commit, FRI layer commit */

/* Goldilocks3::mul, goldilocks_cubic_extension.hpp:1817-1830 (F_p[x]/(x^3 - x - 1)) */
void glo_gl3_mul(uint64_t r[3], const uint64_t a[3], const uint64_t b[3])
{
    uint64_t A = glo_mul(glo_add(a[0], a[1]), glo_add(b[0], b[1]));
    uint64_t B = glo_mul(glo_add(a[0], a[2]), glo_add(b[0], b[2]));
    uint64_t C = glo_mul(glo_add(a[1], a[2]), glo_add(b[1], b[2]));
    uint64_t D = glo_mul(a[0], b[0]), E = glo_mul(a[1], b[1]), F = glo_mul(a[2], b[2]);
    uint64_t G = glo_sub(D, E);
    uint64_t r0 = glo_sub(glo_add(C, G), F);
    uint64_t r1 = glo_sub(glo_sub(glo_sub(glo_add(A, C), E), E), D);
    uint64_t r2 = glo_sub(B, G);
    r[0] = r0; r[1] = r1; r[2] = r2;
}

/* Starks::genProof step 4, starks.cpp:261-284: q_2ns (n_ext x qdim, evaluations on the coset) ->
 * cm4_2ns (n_ext x qdim*qdeg): INTT; qq2[k][p] = qq1[p*n + k] * (7^-n)^p for k < n (rows >= n zero);
 * forward NTT over n_ext. */
void glo_q_extend(uint64_t *cm4, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim, uint64_t qdeg)
{
    uint64_t *qq1 = (uint64_t *)malloc(n_ext * qdim * 8);
    glo_ntt(qq1, q_2ns, n_ext, qdim, 1, 0);                                  /* :261 */
    uint64_t w = qdim * qdeg;
    memset(cm4, 0, n_ext * w * 8);
    uint64_t shift_in = glo_pow(glo_inv(7), n), cur = 1;                      /* :265 */
    for (uint64_t p = 0; p < qdeg; p++) {                                    /* :268-280 */
        for (uint64_t k = 0; k < n; k++)
            for (uint64_t d = 0; d < qdim; d++)
                cm4[k * w + p * qdim + d] = (p * n + k < n_ext) ? glo_mul(qq1[(p * n + k) * qdim + d], cur) : 0;
        cur = glo_mul(cur, shift_in);
    }
    glo_ntt(cm4, cm4, n_ext, w, 0, 0);                                       /* :284 */
    free(qq1);
}

/* FRIProve::prove folding of one step (si > 0), friProve.cpp:36-107: pol (2^pol_bits x 3) -> out
 * (2^next_bits x 3).  For group g: INTT of the nX = 2^(pol_bits-next_bits) elements pol[i*pol2N + g],
 * coefficient i times (shift_inv * w(pol_bits)^-g)^i (polMulAxi :192-200), Horner evaluation at the
 * cubic-extension challenge x (evalPol :201-217). */
void glo_fri_fold(uint64_t *out, const uint64_t *pol, unsigned pol_bits, unsigned next_bits, uint64_t shift_inv, const uint64_t x[3])
{
    uint64_t pol2n = (uint64_t)1 << next_bits, nx = ((uint64_t)1 << pol_bits) / pol2n;
    uint64_t wi = glo_inv(glo_root_of_unity(pol_bits));
#pragma omp parallel for schedule(static)
    for (uint64_t g = 0; g < pol2n; g++) {
        uint64_t *ppar = (uint64_t *)malloc(nx * 3 * 8);
        for (uint64_t i = 0; i < nx; i++) memcpy(ppar + 3 * i, pol + 3 * (i * pol2n + g), 24);
        glo_ntt(ppar, ppar, nx, 3, 1, 0);
        uint64_t sinv = glo_mul(shift_inv, glo_pow(wi, g)), r = 1;
        for (uint64_t i = 0; i < nx; i++) {
            for (int d = 0; d < 3; d++) ppar[3 * i + d] = glo_mul(ppar[3 * i + d], r);
            r = glo_mul(r, sinv);
        }
        uint64_t acc[3] = {ppar[3 * (nx - 1)], ppar[3 * (nx - 1) + 1], ppar[3 * (nx - 1) + 2]};
        for (int64_t i = (int64_t)nx - 2; i >= 0; i--) {
            uint64_t t[3];
            glo_gl3_mul(t, acc, x);
            for (int d = 0; d < 3; d++) acc[d] = glo_add(t[d], ppar[3 * i + d]);
        }
        for (int d = 0; d < 3; d++) out[3 * g + d] = glo_canon(acc[d]);
        free(ppar);
    }
}

/* FRIProve::getTransposed, friProve.cpp:252-270: aux[i*h + j] = pol[j*w + i], w = 2^transpose_bits, 3-word elements */
void glo_fri_transpose(uint64_t *aux, const uint64_t *pol, uint64_t degree, unsigned transpose_bits)
{
    uint64_t w = (uint64_t)1 << transpose_bits, h = degree / w;
    for (uint64_t i = 0; i < w; i++)
        for (uint64_t j = 0; j < h; j++) memcpy(aux + 3 * (i * h + j), pol + 3 * (j * w + i), 24);
}
