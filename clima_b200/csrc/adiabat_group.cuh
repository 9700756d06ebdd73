// adiabat_group.cuh -- the make_profile kernel with GS lanes per column, one lane per gas (included by
// adiabat_profile.cu inside namespace adiabat, after the single-thread building blocks it reuses: Dop,
// dop_integrate, zeroin, sat_pressure, latent_heat, heat_capacity, altitude integration).
//
// Why: the single-thread kernel is bound by the latency of one column's dependent chain, and two thirds of a
// right-hand side is the loop over the gases (saturation pressure = log + exp, heat capacity, latent heat).  Here
// lane g of a group holds gas g (its dry/condensing state, its share of the dry gas, its data), every lane runs
// the integrator itself redundantly on identical numbers -- same instructions, same inputs, so the lanes of a
// group stay bit-identical and their control flow uniform -- and the sums over gases (f_moist, mubar, cp_dry and
// the two sums of Graham et al. eq. 1) are xor-butterflies over the group (the same tree in every lane, and
// a + b = b + a exactly, so every lane gets the same bits).  Reference lines as in adiabat_profile.cu.
#pragma once

template <int GS>
struct Group {
  unsigned mask;  // the lanes of this group within the warp
  int lane;       // 0 .. GS-1 = gas index
  __device__ __forceinline__ double sum(double v) const {
#pragma unroll
    for (int o = GS / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, GS);
    return v;
  }
  __device__ __forceinline__ double bcast(double v, int src) const { return __shfl_sync(mask, v, src, GS); }
  __device__ __forceinline__ int bcast(int v, int src) const { return __shfl_sync(mask, v, src, GS); }
  __device__ __forceinline__ bool any(bool p) const { return (__ballot_sync(mask, p) & mask) != 0u; }
};

// what AdiabatProfileData (:18-71) holds, split over the lanes of a group
template <int GS>
struct GCol {
  const AdiabatData *a;
  Group<GS> g;
  bool act;        // lane < ng
  bool cond;       // CondensingSpeciesType (else DrySpeciesType)
  bool has_sat;
  double f_i_dry;  // this gas's share of the dry gas
  double mass, RH;
  double T_trop;   // of this column (AdiabatOut::T_trop_in) or the object's
  const double *sat;
  int err, nfcn;
};

// mixing_ratios (:548-574): returns this lane's mixing ratio; f_dry by reference (always used by the callers)
template <int GS>
__device__ NI_MISC double g_mixing_ratio(const GCol<GS> &d, double P, double T, double &f_dry) {
  double f = 0.0;
  if (d.act && d.cond) f = fmin(d.RH * sat_pressure(d.sat, T) / P, 1.0);
  const double f_moist = d.g.sum(f);
  f_dry = fmax(1.0 - f_moist, 1.0e-40);
  if (d.act && !d.cond) f = f_dry * d.f_i_dry;
  return f;
}

template <int GS>
__device__ void g_update_f_i_dry(GCol<GS> &d, double P, double f) {  // :529-546
  const double P_dry = d.g.sum((d.act && !d.cond) ? f * P : 0.0);
  d.f_i_dry = (f * P) / P_dry;
}

// root_fcn (:483-513), the component of this lane's gas (the tropopause component T - T_trop is uniform)
template <int GS>
__device__ NI_MISC double g_root_lane(const GCol<GS> &d, double P, double T) {
  double f_dry;
  const double f = g_mixing_ratio(d, P, T, f_dry);
  if (!d.act || d.cond) return 1.0;
  if (!d.has_sat) return DBL_MAX;
  return d.RH * sat_pressure(d.sat, T) - f * P;
}

template <int GS>
struct GProfileRhs {  // right_hand_side (:576-656)
  GCol<GS> *d;
  __device__ NI_RHS void operator()(double P, const double *u, double *du) const {
    GCol<GS> &c = *d;
    const AdiabatData &a = *c.a;
    const double T = u[0], z = u[1];
    const double Rgas_si = RGAS / 1.0e7;
    double f_dry;
    const double f = g_mixing_ratio(c, P, T, f_dry);
    double s_mu = 0.0, s_cpd = 0.0, s_first = 0.0, s_second = 0.0;
    bool bad = false;
    if (c.act) {
      const TPowers tp(T);
      double cp = 0.0;
      bad = !heat_capacity(a, c.g.lane, tp, cp);
      s_mu = f * c.mass;
      if (c.cond) {
        const double L = latent_heat(c.sat, T) * c.mass * 1.0e-7;  // J/mol
        const double beta = L / (Rgas_si * T);
        s_first = f * (cp - Rgas_si * beta + Rgas_si * (beta * beta));
        s_second = beta * f;
      } else {
        s_cpd = c.f_i_dry * cp;
      }
    }
    if (c.g.any(bad)) {
      c.err = 6;
      du[0] = du[1] = 0.0;
      return;
    }
    const double mubar = c.g.sum(s_mu);
    const double cp_dry = DBL_MIN + c.g.sum(s_cpd);
    const double first = c.g.sum(s_first), second = c.g.sum(s_second);
    const double dlnT_dlnP = 1.0 / (f_dry * ((cp_dry * f_dry + first) / (Rgas_si * (f_dry + second))) + second);
    du[0] = dlnT_dlnP * (T / P);
    const double grav = gravity(a.planet_radius, a.planet_mass, z);
    du[1] = -(RGAS * T) / (grav * P * mubar);
    c.nfcn++;
  }
};

template <int GS>
struct GProfileSolout {  // solout_dop (:355-449)
  GCol<GS> *col;
  int ne, j, nz, ng;
  double l0, range, P_surf, Pj;
  double g0_old, gl_old;  // event functions at the previous step: tropopause (uniform), this lane's gas
  int reason, ind_sat;
  double P_root, u_root[2];
  bool z_negative;
  double *T_m, *f_m;

  __device__ NI_SOL double P_edge(int k) const {
    if (k == 0) return P_surf;
    if (k == ne - 1) return col->a->P_top;
    return npow(10.0, l0 + range * k / (ne - 1));
  }
  __device__ void sample(int k, double T, double z, double f) {
    if (z < 0.0 && k < ne - 1) z_negative = true;
    if (k & 1) {
      const int i = k >> 1;
      if (col->g.lane == 0) T_m[i] = T;
      if (col->act) f_m[i + (size_t)nz * col->g.lane] = f;
    }
  }
  __device__ NI_SOL int operator()(const Dop<2> &dop, int nr, double P_old, double P_cur, const double *y) {
    GCol<GS> &d = *col;
    if (d.err) return -10;
    int irtrn = 1;
    const double g0 = y[0] - d.T_trop;
    const double gl = g_root_lane(d, P_cur, y[0]);
    if (nr == 1) {
      g0_old = g0;
      gl_old = gl;
    }
    const bool chg0 = (g0_old < 0.0 && g0 > 0.0) || (g0_old > 0.0 && g0 < 0.0);
    const int chgl = ((gl_old < 0.0 && gl > 0.0) || (gl_old > 0.0 && gl < 0.0)) ? 1 : 0;
    double best = 0.0;
    int ind_root = -1, info = 0;
    if (chg0) {
      const double T_trop = d.T_trop;
      best = zeroin([&](double x) { return dop.contd8(0, x) - T_trop; }, P_old, P_cur, 1.0e-8, info);
      ind_root = 0;
      if (info != 0) d.err = 5;
    }
    if (d.g.any(chgl != 0) && !d.err) {
      for (int i = 0; i < ng; i++) {
        if (!d.g.bcast(chgl, i)) continue;
        const double r = zeroin([&](double x) { return d.g.bcast(g_root_lane(d, x, dop.contd8(0, x)), i); }, P_old,
                                P_cur, 1.0e-8, info);
        if (info != 0) {
          d.err = 5;
          break;
        }
        if (ind_root < 0 || r > best) {  // the highest-pressure root, the first one on a tie (maxloc)
          best = r;
          ind_root = i + 1;
        }
      }
    }
    if (d.err) return -10;
    if (ind_root >= 0) {
      irtrn = -1;
      P_root = best;
      u_root[0] = dop.contd8(0, P_root);
      u_root[1] = dop.contd8(1, P_root);
      if (ind_root == 0) {
        reason = 1;
      } else {
        reason = 2;
        ind_sat = ind_root - 1;
      }
    }
    g0_old = g0;
    gl_old = gl;
    if (j < ne) {
      const double PP = irtrn == -1 ? P_root : P_cur;
      if (Pj <= P_old && Pj >= PP) {
        while (Pj >= PP) {
          const double T = dop.contd8(0, Pj), z = dop.contd8(1, Pj);
          double f_dry;
          const double f = g_mixing_ratio(d, Pj, T, f_dry);
          sample(j, T, z, f);
          j++;
          if (j >= ne) break;
          Pj = P_edge(j);
        }
      }
    }
    return irtrn;
  }
};

// GS lanes per column; `gpw` groups (columns) per warp, the other lanes of the warp leave at once.
template <int GS>
__global__ void __launch_bounds__(ADIABAT_THREADS)
make_profile_group_kernel(const AdiabatData *__restrict__ ap, int ncol, int gpw, const double *__restrict__ T_surf_in,
                          const double *__restrict__ P_i_surf, AdiabatOut o, double *__restrict__ work) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int wl = t & 31, gi = wl / GS;
  if (gi >= gpw) return;
  const int c = (t >> 5) * gpw + gi;
  if (c >= ncol) return;
  const AdiabatData &a = *ap;
  const int ng = a.ng, nz = a.nz, ne = 2 * nz + 1;
  GCol<GS> col;
  col.a = ap;
  col.g.mask = (GS == 32 ? 0xffffffffu : ((1u << GS) - 1u)) << (gi * GS);
  col.g.lane = wl % GS;
  const int lane = col.g.lane;
  col.act = lane < ng;
  col.err = 0;
  col.nfcn = 0;
  const int gl = col.act ? lane : 0;  // a valid gas index for the loads of idle lanes
  col.mass = a.mass[gl];
  col.RH = a.RH[gl];
  col.has_sat = col.act && a.has_sat[gl];
  col.sat = a.sat[gl];
  col.cond = false;
  col.f_i_dry = 0.0;
  const double T_surf = T_surf_in[c];
  const double Pin = col.act ? P_i_surf[(size_t)ng * c + lane] : 0.0;
  double *T_m = o.T + (size_t)nz * c, *P_m = o.P + (size_t)nz * c, *z_m = o.z + (size_t)nz * c,
         *dz_m = o.dz + (size_t)nz * c, *f_m = o.f_i + (size_t)nz * ng * c;
  int status = 0;
  if (lane == 0) {
    o.P_trop[c] = -1.0;
    o.P_surf[c] = 0.0;
    o.nfcn[c] = 0;
  }
  // input checks (:124-139)
  const double T_trop = o.T_trop_in ? o.T_trop_in[c] : a.T_trop;
  col.T_trop = T_trop;
  if (T_surf < T_trop) status = 1;
  if (!status && col.g.any(Pin < 0.0)) status = 2;
  if (!status && T_trop < 0.0) status = 10;
  if (status) {
    if (lane == 0) o.status[c] = status;
    return;
  }
  // which gases condense at the surface (:199-224)
  const double grav0 = gravity(a.planet_radius, a.planet_mass, 0.0);
  double f_cur = 0.0;
  if (col.act) {
    const double P_sat = col.has_sat ? col.RH * sat_pressure(col.sat, T_surf) : DBL_MAX;
    double N_surf = 0.0;
    if (Pin > P_sat) {
      f_cur = P_sat;
      N_surf = (Pin - P_sat) / (col.mass * grav0);
      col.cond = true;
    } else {
      f_cur = Pin;
    }
    o.N_surface[(size_t)ng * c + lane] = N_surf;
  }
  const double P_surf = col.g.sum(f_cur);
  if (lane == 0) o.P_surf[c] = P_surf;
  if (a.P_top > P_surf) {
    if (lane == 0) o.status[c] = 3;
    return;
  }
  f_cur /= P_surf;
  g_update_f_i_dry(col, P_surf, f_cur);
  if (col.act) o.f_i_surf[(size_t)ng * c + lane] = f_cur;
  const double mubar_surf = col.g.sum(f_cur * col.mass);

  GProfileSolout<GS> so;
  so.col = &col;
  so.ne = ne;
  so.j = 1;
  so.nz = nz;
  so.ng = ng;
  so.l0 = nlog10(P_surf);
  so.range = nlog10(a.P_top) - so.l0;
  so.P_surf = P_surf;
  so.Pj = so.P_edge(1);
  so.reason = 0;
  so.z_negative = false;
  so.T_m = T_m;
  so.f_m = f_m;
  if (lane == 0)
    for (int i = 0; i < nz; i++) P_m[i] = so.P_edge(2 * i + 1);
  {
    Dop<2> dop;
    GProfileRhs<GS> rhs{&col};
    double u[2] = {T_surf, 0.0}, Pn = P_surf;
    for (;;) {  // integrate (:274-353)
      so.reason = 0;
      const int idid = dop_integrate<2>(dop, Pn, u, a.P_top, a.rtol, a.atol, rhs, so);
      if (col.err) {
        status = col.err;
        break;
      }
      if (idid < 0) {
        status = 4;
        break;
      }
      if (so.reason != 2) break;
      Pn = so.P_root;
      u[0] = so.u_root[0];
      u[1] = so.u_root[1];
      double f_dry;
      f_cur = g_mixing_ratio(col, Pn, u[0], f_dry);
      if (lane == so.ind_sat) col.cond = true;
      g_update_f_i_dry(col, Pn, f_cur);
    }
  }
  if (!status && so.reason == 1) {  // isothermal stratosphere (:331-351)
    double f_dry;
    f_cur = g_mixing_ratio(col, so.P_root, so.u_root[0], f_dry);
    if (lane == 0) o.P_trop[c] = so.P_root;
    const double mubar = col.g.sum(f_cur * col.mass);
    for (int k = so.j; k < ne; k++)
      so.sample(k, T_trop,
                altitude_vs_P(so.P_edge(k), T_trop, mubar, so.P_root, so.u_root[1], a.planet_mass, a.planet_radius),
                f_cur);
  }
  if (!status && so.z_negative) status = 7;
  if (lane == 0) o.nfcn[c] = col.nfcn;
  if (status) {
    if (lane == 0) o.status[c] = status;
    return;
  }
  __syncwarp(col.g.mask);  // T_m / P_m of lane 0 and every lane's f_m are visible to the group

  // compute_altitude (clima_adiabat_altitude.f90:41-193): the tables by the group, the integration by lane 0
  double *lp = work + (size_t)c * (3 * (nz + 1) + 2 * ne), *Tg = lp + nz + 1, *mg = Tg + nz + 1, *P_e2 = mg + nz + 1,
         *z_e = P_e2 + ne;
  for (int i = 0; i < nz; i++) {
    const double m = col.g.sum(col.act ? f_m[i + (size_t)nz * lane] * col.mass : 0.0);
    if (lane == 0) mg[i + 1] = m;
  }
  if (lane == 0) {
    lp[0] = nlog10(P_surf);
    Tg[0] = T_surf;
    mg[0] = mubar_surf;
    P_e2[0] = P_surf;
    for (int i = 0; i < nz; i++) {
      lp[i + 1] = nlog10(P_m[i]);
      Tg[i + 1] = T_m[i];
      P_e2[2 * i + 1] = P_m[i];
      if (i < nz - 1) P_e2[2 * i + 2] = sqrt(P_m[i] * P_m[i + 1]);
    }
    P_e2[2 * nz] = a.P_top;
    status = altitude_profile(a, nz, P_surf, lp, Tg, mg, P_e2, z_e);
    if (!status)
      for (int i = 0; i < nz; i++) {
        z_m[i] = z_e[2 * i + 1];
        dz_m[i] = z_e[2 * i + 2] - z_e[2 * i];
      }
    o.status[c] = status;
  }
  status = col.g.bcast(status, 0);
  if (status) return;
  __syncwarp(col.g.mask);

  // densities (clima_adiabat.f90:446-449) and copy_atm_to_radiative_grid (:729-773): lane g writes gas g
  if (o.T_r) {
    const bool dbl = a.double_radiative_grid != 0;
    const int nzr = dbl ? 2 * nz + 2 : nz;
    double *T_r = o.T_r + (size_t)nzr * c, *P_r = o.P_r + (size_t)nzr * c, *dz_r = o.dz_r + (size_t)nzr * c,
           *dens_r = o.dens_r + (size_t)nzr * ng * c;
    for (int i = 0; i < nz; i++) {
      const double density = P_m[i] / (K_BOLTZ * T_m[i]);
      const int reps = dbl ? (i == nz - 1 ? 4 : 2) : 1;
      const int j0 = dbl ? 2 * i : i;
      const double dzr = dbl ? 0.5 * dz_m[i] : dz_m[i];
      const double dens = col.act ? f_m[i + (size_t)nz * lane] * density : 0.0;
      for (int r = 0; r < reps; r++) {
        if (lane == 0) {
          T_r[j0 + r] = T_m[i];
          P_r[j0 + r] = P_m[i] / 1.0e6;
          dz_r[j0 + r] = dzr;
        }
        if (col.act) dens_r[j0 + r + (size_t)nzr * lane] = dens;
      }
    }
    if (lane == 0) o.Tsurf_r[c] = T_surf;
    write_particles(a, o, c, P_m, lane, GS);
  }
}
