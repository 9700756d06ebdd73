/* tests/c_driver.c -- a caller of the C ABI that is not Python: plain C against include/clima_rt.h,
 * the closest thing to the Fortran `bind(c)` caller this box can compile (no Fortran compiler).
 *
 *   c_driver case.crtb
 *
 * reads a case file written by clima_b200/blob.py (tables, params, a batch of columns, the oracle's
 * integrated fluxes and optionally a frozen-opacity Jacobian case), then
 *   clima_rt_create -> clima_rt_radiate_batch -> compare with the expected fluxes (1e-9 relative)
 *   -> [clima_rt_radiate_batch on column 0 + clima_rt_jacobian_batch -> compare] -> clima_rt_destroy
 * following the reference's conventions: scalars by reference, err[1025], "" = success
 * (clima/fortran/clima_c_api.f90:5,57-69).  Exit code 0 and a line "PASS" on success. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "clima_rt.h"

static FILE *F;
static void die(const char *m) { fprintf(stderr, "c_driver: %s\n", m); exit(2); }
static void *rec(int kind, int *n) {
  int32_t hd[2];
  if (fread(hd, 4, 2, F) != 2) die("short file");
  if (hd[0] != kind) die("record kind mismatch");
  *n = hd[1];
  const size_t sz = kind == 0 ? 4 : 8;
  void *p = malloc((size_t)(hd[1] > 0 ? hd[1] : 1) * sz);
  if (hd[1] > 0 && fread(p, sz, (size_t)hd[1], F) != (size_t)hd[1]) die("short record");
  return p;
}
static int32_t *ri(int *n) { return (int32_t *)rec(0, n); }
static double *rd(int *n) { return (double *)rec(1, n); }
static double relerr(const double *a, const double *b, int n) {
  double s = 0.0, d = 0.0;
  for (int i = 0; i < n; i++) { if (fabs(b[i]) > s) s = fabs(b[i]); if (fabs(a[i] - b[i]) > d) d = fabs(a[i] - b[i]); }
  return s > 0.0 ? d / s : d;
}

int main(int argc, char **argv) {
  if (argc < 2) die("usage: c_driver case.crtb");
  F = fopen(argv[1], "rb");
  if (!F) die("cannot open the case file");
  char magic[8];
  if (fread(magic, 1, 8, F) != 8 || memcmp(magic, "CRTB1\0\0\0", 8) != 0) die("bad magic");
  int n;
  clima_rt_tables t;
  memset(&t, 0, sizeof t);
  int32_t *h = ri(&n);
  if (n != 16) die("bad header");
  t.nw = h[0]; t.ng = h[1]; t.np = h[2]; t.nk = h[3]; t.ngauss = h[4]; t.ncia = h[5]; t.nray = h[6]; t.npxs = h[7];
  t.has_cont = h[8]; t.cont_LH2O = h[9]; t.cont_ntemp = h[10]; t.npart = h[11];
  t.ir_ind_start = h[12]; t.ir_ind_end = h[13]; t.sol_ind_start = h[14]; t.sol_ind_end = h[15];
  t.wavl = rd(&n); t.weights = rd(&n);
  t.k_sp_ind = ri(&n); t.k_npress = ri(&n); t.k_ntemp = ri(&n);
  const double **kP = malloc(sizeof(double *) * (size_t)(t.nk + 1)), **kT = malloc(sizeof(double *) * (size_t)(t.nk + 1)),
               **kk = malloc(sizeof(double *) * (size_t)(t.nk + 1));
  for (int i = 0; i < t.nk; i++) { kP[i] = rd(&n); kT[i] = rd(&n); kk[i] = rd(&n); }
  t.k_log10P = kP; t.k_temp = kT; t.k_log10k = kk;
  t.cia_sp_ind = ri(&n); t.cia_dim = ri(&n); t.cia_ntemp = ri(&n);
  const double **ct = malloc(sizeof(double *) * (size_t)(t.ncia + 1)), **cd = malloc(sizeof(double *) * (size_t)(t.ncia + 1));
  for (int i = 0; i < t.ncia; i++) { ct[i] = rd(&n); cd[i] = rd(&n); }
  t.cia_temp = ct; t.cia_data = cd;
  t.ray_sp_ind = ri(&n); t.ray_xs = rd(&n);
  t.pxs_sp_ind = ri(&n); t.pxs_dim = ri(&n); t.pxs_ntemp = ri(&n);
  const double **pt = malloc(sizeof(double *) * (size_t)(t.npxs + 1)), **pd = malloc(sizeof(double *) * (size_t)(t.npxs + 1));
  for (int i = 0; i < t.npxs; i++) { pt[i] = rd(&n); pd[i] = rd(&n); }
  t.pxs_temp = pt; t.pxs_data = pd;
  t.cont_temp = rd(&n); t.cont_log10_H2O = rd(&n); t.cont_log10_foreign = rd(&n);
  t.part_p_ind = ri(&n); t.part_nrad = ri(&n);
  const double **pr_ = malloc(sizeof(double *) * (size_t)(t.npart + 1)), **pw = malloc(sizeof(double *) * (size_t)(t.npart + 1)),
               **pq = malloc(sizeof(double *) * (size_t)(t.npart + 1)), **pg = malloc(sizeof(double *) * (size_t)(t.npart + 1));
  for (int i = 0; i < t.npart; i++) { pr_[i] = rd(&n); pw[i] = rd(&n); pq[i] = rd(&n); pg[i] = rd(&n); }
  t.part_radii = pr_; t.part_w0 = pw; t.part_qext = pq; t.part_gt = pg;
  t.photons_sol = rd(&n);

  clima_rt_params p;
  memset(&p, 0, sizeof p);
  double *ps = rd(&n);
  p.diurnal_fac = ps[0]; p.photon_scale_factor = ps[1]; p.ir_tau_min = ps[2];
  int32_t *pi = ri(&n);
  p.has_hard_surface = pi[0]; p.nzenith = pi[1];
  p.zenith_u = rd(&n); p.zenith_weights = rd(&n); p.surface_albedo = rd(&n); p.surface_emissivity = rd(&n);

  int32_t *cd4 = ri(&n);
  const int32_t ncol = cd4[0], nz = cd4[1], ng = cd4[2], np = cd4[3];
  double *Ts = rd(&n), *T = rd(&n), *P = rd(&n), *dens = rd(&n), *dz = rd(&n), *pdn = NULL, *rad = NULL;
  if (np > 0) { pdn = rd(&n); rad = rd(&n); }
  double *gold[7];
  for (int k = 0; k < 7; k++) gold[k] = rd(&n);

  char err[CLIMA_RT_ERR_LEN + 1];
  void *handle = NULL;
  const int32_t dev = 0, one = 1, zero = 0;
  clima_rt_create(&t, &p, &dev, &handle, err);
  if (err[0]) { fprintf(stderr, "clima_rt_create: %s\n", err); return 1; }

  const size_t L = (size_t)nz + 1;
  clima_rt_out o;
  memset(&o, 0, sizeof o);
  double *res[7];
  for (int k = 0; k < 7; k++) res[k] = calloc((k < 5 ? L : 1) * (size_t)ncol, 8);
  o.fup_n_ir = res[0]; o.fdn_n_ir = res[1]; o.fup_n_sol = res[2]; o.fdn_n_sol = res[3]; o.f_total = res[4];
  o.ISR = res[5]; o.OLR = res[6];
  clima_rt_radiate_batch(handle, &ncol, &nz, &ng, &np, Ts, T, P, dens, dz, pdn, rad, &one, &one, &o, err);
  if (err[0]) { fprintf(stderr, "clima_rt_radiate_batch: %s\n", err); return 1; }
  static const char *names[7] = {"fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR"};
  int bad = 0;
  double scale = 0.0;
  for (int k = 0; k < 4; k++) for (size_t i = 0; i < L * (size_t)ncol; i++) if (fabs(gold[k][i]) > scale) scale = fabs(gold[k][i]);
  for (int k = 0; k < 7; k++) {
    double e = relerr(res[k], gold[k], (int)((k < 5 ? L : 1) * (size_t)ncol));
    if (k == 4) {  /* f_total: a difference of four large numbers, against the flux scale */
      double d = 0.0;
      for (size_t i = 0; i < L * (size_t)ncol; i++) if (fabs(res[4][i] - gold[4][i]) > d) d = fabs(res[4][i] - gold[4][i]);
      e = d / scale;
    }
    printf("%-10s max rel err %.3e\n", names[k], e);
    if (!(e <= 1.0e-9)) bad = 1;
  }

  /* error convention: a wrong dimension must come back as a message, not a crash */
  const int32_t ng_bad = ng - 1;
  clima_rt_radiate_batch(handle, &ncol, &nz, &ng_bad, &np, Ts, T, P, dens, dz, pdn, rad, &one, &one, &o, err);
  if (!strstr(err, "densities")) { fprintf(stderr, "expected the \"densities\" dimension error, got \"%s\"\n", err); bad = 1; }

  int32_t *jh = ri(&n);
  if (jh[0]) {
    const int32_t ncj = jh[1], nzm = jh[2];
    double *Tsj = rd(&n), *Tj = rd(&n), *fj = rd(&n);
    /* the objective evaluation at the base state (column 0) retains opacity + solar fluxes */
    clima_rt_radiate_batch(handle, &one, &nz, &ng, &np, Ts, T, P, dens, dz, pdn, rad, &one, &one, &o, err);
    if (err[0]) { fprintf(stderr, "base column: %s\n", err); return 1; }
    double *f = calloc((size_t)(nzm + 1) * (size_t)ncj, 8);
    clima_rt_jacobian_batch(handle, &ncj, &nzm, Tsj, Tj, f, err);
    if (err[0]) { fprintf(stderr, "clima_rt_jacobian_batch: %s\n", err); return 1; }
    double d = 0.0;
    for (size_t i = 0; i < (size_t)(nzm + 1) * (size_t)ncj; i++) if (fabs(f[i] - fj[i]) > d) d = fabs(f[i] - fj[i]);
    printf("jacobian   max err / flux scale %.3e\n", d / scale);
    if (!(d / scale <= 1.0e-9)) bad = 1;
  }
  (void)zero;
  clima_rt_destroy(handle);
  puts(bad ? "FAIL" : "PASS");
  return bad;
}
