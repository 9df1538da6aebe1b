/* ani_oracle.c -- CPU (fp64, OpenMP) restatement of lammps-ani's per-timestep ANI energy/force path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (lammps_ani_b200/, include/) links, imports or
 * calls this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs do, and only as the checker or the CPU baseline.
 *
 * PARITY STATUS: "parity unpinned" against the reference's absolute golden numbers.  The arithmetic
 * of this path lives in the third-party torchani submodule (aiqm/torchani pinned at
 * 4f56c736caea0221f5fefa0e61e5bb5fc2fd1142, /root/reference/.SUBMODULES.json: missing) and every
 * golden vector in the reference tree (src/ani_csrc/test_model.cpp:84-137,
 * tests/lammps-unittest/ * / *.yaml) is for ANI-2x, whose weights are not available offline.  The
 * oracle is therefore pinned by (1) a second, independent restatement that mirrors the pyaev
 * pair/triple formulation in torch fp64 with autograd (oracle/pyaev_torch.py), (2) central finite
 * differences of E for forces and virial, (3) the in-tree ANI-1xnr weights for real-parameter runs.
 *
 * What it follows:
 *   step graph, ghost/padding/sign/unit conventions ... /root/reference/models/lammps_ani.py:131-257
 *   (E sums LOCAL atoms only: ghosts are species -1 padding, lammps_ani.py:228-230, ani.cpp:226-228;
 *    force = -dE/dx for local AND ghost atoms, lammps_ani.py:203-214;
 *    virial = -sym(sum_pairs dE/d(diff) (x) diff), lammps_ani.py:197-200,215)
 *   AEV formulas: published ANI descriptor (Smith et al. 2017) as stated in SURVEY.md 8.1:
 *    radial  G[s_j*nR+k]            += 0.25 exp(-EtaR (r-ShfR_k)^2) fc(r;Rcr)
 *    angular G[S*nR + pair(s_j,s_k)*nA*nZ + a*nZ + z]
 *                                   += 2 ((1+cos(th-ShfZ_z))/2)^Zeta exp(-EtaA((rj+rk)/2-ShfA_a)^2) fc(rj;Rca) fc(rk;Rca)
 *    th = acos(0.95 cos_jik), fc(r;Rc) = 0.5 cos(pi r/Rc)+0.5, pair(s,t) upper-triangular row-major.
 *   MLP: CELU(alpha=0.1) hidden layers, linear output, mean over the first `num_models` members
 *    (select_models, lammps_ani.py:332-343), plus self atomic energies.
 * Units: Hartree and Angstrom; the caller multiplies by 627.5094738898777 (ani.h:9, ani.cpp:161-162).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXS 16
#define MAXL 8

typedef struct {
  int S, nR, nA, nZ, M, maxl, use_rep;
  /* RepulsionXTB pair term (models/lammps_ani.py:186-192,301-330; torchani potentials.RepulsionXTB, not in tree) */
  double rep_cutoff; int rep_fn; const double *rep_y, *rep_sa, *rep_k;
  double Rcr, Rca, EtaR, EtaA, Zeta;
  double *ShfR, *ShfA, *ShfZ, *sae;
  int nlayers[MAXS];
  int dims[MAXS][MAXL + 1];
  float** W; /* [M*S*MAXL] */
  float** b;
  int aev_len, rad_len, ang_sub;
  int triu[MAXS][MAXS];
  unsigned char* blob;
} omodel;

static int rd_i32(const unsigned char* p) { int v; memcpy(&v, p, 4); return v; }

omodel* ani_oracle_model_load(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) return NULL;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  unsigned char* buf = (unsigned char*)malloc(n);
  if (fread(buf, 1, n, f) != (size_t)n) { fclose(f); free(buf); return NULL; }
  fclose(f);
  if (memcmp(buf, "ANIW0001", 8) != 0) { free(buf); return NULL; }
  omodel* m = (omodel*)calloc(1, sizeof(omodel));
  m->blob = buf;
  long off = 8;
  m->S = rd_i32(buf + off); m->nR = rd_i32(buf + off + 4); m->nA = rd_i32(buf + off + 8);
  m->nZ = rd_i32(buf + off + 12); m->M = rd_i32(buf + off + 16); m->maxl = rd_i32(buf + off + 20);
  m->use_rep = rd_i32(buf + off + 24);
  off += 32;
  double c[5]; memcpy(c, buf + off, 40); off += 40;
  m->Rcr = c[0]; m->Rca = c[1]; m->EtaR = c[2]; m->EtaA = c[3]; m->Zeta = c[4];
  m->ShfR = (double*)malloc(8 * m->nR); memcpy(m->ShfR, buf + off, 8 * m->nR); off += 8 * m->nR;
  m->ShfA = (double*)malloc(8 * m->nA); memcpy(m->ShfA, buf + off, 8 * m->nA); off += 8 * m->nA;
  m->ShfZ = (double*)malloc(8 * m->nZ); memcpy(m->ShfZ, buf + off, 8 * m->nZ); off += 8 * m->nZ;
  off += 4 * m->S; /* symbols */
  m->sae = (double*)malloc(8 * m->S); memcpy(m->sae, buf + off, 8 * m->S); off += 8 * m->S;
  for (int s = 0; s < m->S; s++) {
    m->nlayers[s] = rd_i32(buf + off); off += 4;
    for (int l = 0; l <= m->nlayers[s]; l++) { m->dims[s][l] = rd_i32(buf + off); off += 4; }
  }
  m->W = (float**)calloc((size_t)m->M * m->S * MAXL, sizeof(float*));
  m->b = (float**)calloc((size_t)m->M * m->S * MAXL, sizeof(float*));
  for (int mm = 0; mm < m->M; mm++)
    for (int s = 0; s < m->S; s++)
      for (int l = 0; l < m->nlayers[s]; l++) {
        size_t idx = ((size_t)mm * m->S + s) * MAXL + l;
        m->W[idx] = (float*)(buf + off); off += 4L * m->dims[s][l] * m->dims[s][l + 1];
        m->b[idx] = (float*)(buf + off); off += 4L * m->dims[s][l + 1];
      }
  if (m->use_rep) {
    memcpy(&m->rep_cutoff, buf + off, 8); m->rep_fn = rd_i32(buf + off + 8); off += 16;
    m->rep_y = (const double*)(buf + off); off += 8L * m->S * m->S;
    m->rep_sa = (const double*)(buf + off); off += 8L * m->S * m->S;
    m->rep_k = (const double*)(buf + off); off += 8L * m->S * m->S;
  }
  if (off != n) { fprintf(stderr, "ani_oracle: trailing bytes in %s\n", path); }
  m->rad_len = m->S * m->nR;
  m->ang_sub = m->nA * m->nZ;
  m->aev_len = m->rad_len + (m->S * (m->S + 1) / 2) * m->ang_sub;
  int p = 0;
  for (int s = 0; s < m->S; s++)
    for (int t = s; t < m->S; t++) { m->triu[s][t] = p; m->triu[t][s] = p; p++; }
  return m;
}

void ani_oracle_model_free(omodel* m) {
  if (!m) return;
  free(m->ShfR); free(m->ShfA); free(m->ShfZ); free(m->sae); free(m->W); free(m->b); free(m->blob); free(m);
}

int ani_oracle_aev_length(const omodel* m) { return m->aev_len; }
int ani_oracle_num_models(const omodel* m) { return m->M; }
int ani_oracle_num_species(const omodel* m) { return m->S; }

static inline double celu(double x) { return x > 0 ? x : 0.1 * (exp(x / 0.1) - 1.0); }
/* Backward of CELU as the reference's autograd evaluates it.  torch's derivative formula for celu is
 * elu_backward(grad, alpha, 1, 1.0/alpha.toFloat(), ...): the input scale is 1/(float)0.1 =
 * 9.99999985098839 while the forward uses exactly 1/alpha, so for x <= 0
 *   dCELU = 0.1 * k * exp(k x),  k = 1/(double)(float)0.1
 * (a 1.5e-8 relative deviation from exp(x/0.1), visible at the fp64 tolerance of the reference's goldens,
 * 9e-9 .. 1e-13).  g_celu_exact = 1 switches to the mathematically exact derivative. */
static int g_celu_exact = 0;
void ani_oracle_set_celu_exact(int on) { g_celu_exact = on; }
static inline double dcelu(double x) {
  if (x > 0) return 1.0;
  if (g_celu_exact) return exp(x / 0.1);
  const double k = 1.0 / (double)0.1f;
  return 0.1 * k * exp(k * x);
}

typedef struct { double d[3]; double r; int j; int s; } onbr;

/* AEV row of one centre from its neighbour set (already holding displacement d = x_j - x_i). */
static void aev_row(const omodel* m, int radial_filter, const onbr* nb, int nn, double* G) {
  const double PI = 3.14159265358979323846;
  memset(G, 0, sizeof(double) * m->aev_len);
  for (int a = 0; a < nn; a++) {
    double r = nb[a].r;
    if (radial_filter && r > m->Rcr) continue;
    double fc = 0.5 * cos(PI * r / m->Rcr) + 0.5;
    double* g = G + nb[a].s * m->nR;
    for (int k = 0; k < m->nR; k++) {
      double dr = r - m->ShfR[k];
      g[k] += 0.25 * exp(-m->EtaR * dr * dr) * fc;
    }
  }
  for (int a = 0; a < nn; a++) {
    if (nb[a].r > m->Rca) continue;
    double fca = 0.5 * cos(PI * nb[a].r / m->Rca) + 0.5;
    for (int b = a + 1; b < nn; b++) {
      if (nb[b].r > m->Rca) continue;
      double fcb = 0.5 * cos(PI * nb[b].r / m->Rca) + 0.5;
      double dot = nb[a].d[0] * nb[b].d[0] + nb[a].d[1] * nb[b].d[1] + nb[a].d[2] * nb[b].d[2];
      double den = nb[a].r * nb[b].r;
      if (den < 1e-10) den = 1e-10;
      double th = acos(0.95 * dot / den);
      double ravg = 0.5 * (nb[a].r + nb[b].r);
      double* g = G + m->rad_len + m->triu[nb[a].s][nb[b].s] * m->ang_sub;
      for (int ia = 0; ia < m->nA; ia++) {
        double dr = ravg - m->ShfA[ia];
        double f2 = exp(-m->EtaA * dr * dr);
        for (int iz = 0; iz < m->nZ; iz++) {
          double f1 = pow(0.5 * (1.0 + cos(th - m->ShfZ[iz])), m->Zeta);
          g[ia * m->nZ + iz] += 2.0 * f1 * f2 * fca * fcb;
        }
      }
    }
  }
}

/* Given A = dE/dG for one centre, accumulate gvec[a] = dE_i/d(d_a) for each neighbour. */
static void aev_row_backward(const omodel* m, int radial_filter, const onbr* nb, int nn, const double* A,
                             double (*gvec)[3]) {
  const double PI = 3.14159265358979323846;
  for (int a = 0; a < nn; a++) gvec[a][0] = gvec[a][1] = gvec[a][2] = 0.0;
  for (int a = 0; a < nn; a++) {
    double r = nb[a].r;
    if (radial_filter && r > m->Rcr) continue;
    double fc = 0.5 * cos(PI * r / m->Rcr) + 0.5;
    double dfc = -0.5 * PI / m->Rcr * sin(PI * r / m->Rcr);
    const double* g = A + nb[a].s * m->nR;
    double dEdr = 0.0;
    for (int k = 0; k < m->nR; k++) {
      double dr = r - m->ShfR[k];
      double e = 0.25 * exp(-m->EtaR * dr * dr);
      dEdr += g[k] * e * (dfc - 2.0 * m->EtaR * dr * fc);
    }
    for (int c = 0; c < 3; c++) gvec[a][c] += dEdr * nb[a].d[c] / r;
  }
  for (int a = 0; a < nn; a++) {
    if (nb[a].r > m->Rca) continue;
    double ra = nb[a].r;
    double fca = 0.5 * cos(PI * ra / m->Rca) + 0.5;
    double dfca = -0.5 * PI / m->Rca * sin(PI * ra / m->Rca);
    for (int b = a + 1; b < nn; b++) {
      if (nb[b].r > m->Rca) continue;
      double rb = nb[b].r;
      double fcb = 0.5 * cos(PI * rb / m->Rca) + 0.5;
      double dfcb = -0.5 * PI / m->Rca * sin(PI * rb / m->Rca);
      double dot = nb[a].d[0] * nb[b].d[0] + nb[a].d[1] * nb[b].d[1] + nb[a].d[2] * nb[b].d[2];
      double den = ra * rb;
      int clamped = 0;
      if (den < 1e-10) { den = 1e-10; clamped = 1; }
      double cosv = dot / den;
      double c = 0.95 * cosv;
      double sn = sqrt(1.0 - c * c);
      double th = acos(c);
      double ravg = 0.5 * (ra + rb);
      const double* g = A + m->rad_len + m->triu[nb[a].s][nb[b].s] * m->ang_sub;
      double dT_dra = 0, dT_drb = 0, dT_dcos = 0;
      for (int ia = 0; ia < m->nA; ia++) {
        double dr = ravg - m->ShfA[ia];
        double f2 = exp(-m->EtaA * dr * dr);
        double df2 = -2.0 * m->EtaA * dr * f2; /* d f2 / d ravg */
        for (int iz = 0; iz < m->nZ; iz++) {
          double u = 0.5 * (1.0 + cos(th - m->ShfZ[iz]));
          double f1 = pow(u, m->Zeta);
          /* d u/d c with th = acos(c): d cos(th - s)/dc = sin(th - s)/sin(th) */
          double du_dc = 0.5 * sin(th - m->ShfZ[iz]) / sn;
          double df1_dc = (u > 0) ? m->Zeta * pow(u, m->Zeta - 1.0) * du_dc : 0.0;
          double w = g[ia * m->nZ + iz];
          dT_dra += w * 2.0 * f1 * fcb * (0.5 * df2 * fca + f2 * dfca);
          dT_drb += w * 2.0 * f1 * fca * (0.5 * df2 * fcb + f2 * dfcb);
          dT_dcos += w * 2.0 * f2 * fca * fcb * df1_dc * 0.95;
        }
      }
      for (int cc = 0; cc < 3; cc++) {
        double ua = nb[a].d[cc] / ra, ub = nb[b].d[cc] / rb;
        double dcos_da = clamped ? nb[b].d[cc] / den : (ub - cosv * ua) / ra;
        double dcos_db = clamped ? nb[a].d[cc] / den : (ua - cosv * ub) / rb;
        gvec[a][cc] += dT_dra * ua + dT_dcos * dcos_da;
        gvec[b][cc] += dT_drb * ub + dT_dcos * dcos_db;
      }
    }
  }
}

/* Per-atom ensemble MLP: returns mean_m e_m and fills dG = d(mean_m e_m)/dG.
 * work must hold (2*nl+2)*maxw doubles.                                                   */
static double mlp_atom(const omodel* m, int nm, int s, const double* G, double* dG, double* work, int maxw) {
  int nl = m->nlayers[s];
  const int* dims = m->dims[s];
  double* z = work;                            /* [nl][maxw] pre-activations */
  double* h = work + (size_t)nl * maxw;        /* [nl][maxw] activations     */
  double* gcur = work + (size_t)2 * nl * maxw; /* [maxw] */
  double* gnext = gcur + maxw;                 /* [maxw] */
  double etot = 0.0;
  for (int k = 0; k < dims[0]; k++) dG[k] = 0.0;
  for (int mm = 0; mm < nm; mm++) {
    const double* in = G;
    for (int l = 0; l < nl; l++) {
      const float* W = m->W[((size_t)mm * m->S + s) * MAXL + l];
      const float* b = m->b[((size_t)mm * m->S + s) * MAXL + l];
      int ni = dims[l], no = dims[l + 1];
      double* zl = z + (size_t)l * maxw;
      double* hl = h + (size_t)l * maxw;
      for (int o = 0; o < no; o++) {
        const float* w = W + (size_t)o * ni;
        double acc = 0.0;
        for (int k = 0; k < ni; k++) acc += (double)w[k] * in[k];
        zl[o] = acc + (double)b[o];
      }
      if (l < nl - 1) {
        for (int o = 0; o < no; o++) hl[o] = celu(zl[o]);
        in = hl;
      }
    }
    etot += z[(size_t)(nl - 1) * maxw];
    /* backward: gcur = dE/dz_l */
    gcur[0] = 1.0;
    for (int l = nl - 1; l >= 0; l--) {
      const float* W = m->W[((size_t)mm * m->S + s) * MAXL + l];
      int ni = dims[l], no = dims[l + 1];
      if (l == 0) {
        for (int o = 0; o < no; o++) {
          const float* w = W + (size_t)o * ni;
          double go = gcur[o];
          for (int k = 0; k < ni; k++) dG[k] += go * (double)w[k];
        }
      } else {
        for (int k = 0; k < ni; k++) gnext[k] = 0.0;
        for (int o = 0; o < no; o++) {
          const float* w = W + (size_t)o * ni;
          double go = gcur[o];
          for (int k = 0; k < ni; k++) gnext[k] += go * (double)w[k];
        }
        const double* zprev = z + (size_t)(l - 1) * maxw;
        for (int k = 0; k < ni; k++) gnext[k] *= dcelu(zprev[k]);
        double* t = gcur; gcur = gnext; gnext = t;
      }
    }
  }
  double inv = 1.0 / nm;
  for (int k = 0; k < dims[0]; k++) dG[k] *= inv;
  return etot * inv;
}

/* RepulsionXTB pair energy (Hartree) and its derivative with respect to r (Angstrom):
 *   e(r) = y_ab / r_b * exp(-sqrt_alpha_ab * r_b^k_ab) * fc(r),  r_b = r in Bohr   (GFN2-xTB repulsion; torchani potentials.RepulsionXTB)
 *   fc: 1 = exp(1 - 1 / (1 - (r/Rc)^2)) ("smooth", order 2);  0 = 0.5 cos(pi r / Rc) + 0.5.
 * Every list ENTRY (centre i local, neighbour j) carries half of the pair: a local-local pair appears in both rows, a pair with
 * a ghost only once -- torchani halves pairs that touch a ghost (`ghost_flags`), the other half belongs to the neighbouring
 * subdomain / periodic image.  Restated from the published functional form: the torchani source is not in the tree and no model
 * file with `rep_calc` is available, so this term is PARITY UNPINNED. */
#define ANGSTROM_TO_BOHR 1.8897261258369282
static double rep_pair(const omodel* m, int sa_, int sb_, double r, double* de_dr) {
  const double Rc = m->rep_cutoff;
  if (r > Rc || r <= 0.0) { *de_dr = 0.0; return 0.0; }
  const double y = m->rep_y[sa_ * m->S + sb_], sa = m->rep_sa[sa_ * m->S + sb_], k = m->rep_k[sa_ * m->S + sb_];
  const double rb = r * ANGSTROM_TO_BOHR;
  const double p = pow(rb, k);
  const double base = y / rb * exp(-sa * p);
  const double dbase = base * (-1.0 / r - sa * k * p / r);
  double fc, dfc;
  if (m->rep_fn == 1) {
    const double t = (r / Rc) * (r / Rc);
    if (t >= 1.0) { *de_dr = 0.0; return 0.0; }
    fc = exp(1.0 - 1.0 / (1.0 - t));
    dfc = fc * (-1.0 / ((1.0 - t) * (1.0 - t))) * (2.0 * r / (Rc * Rc));
  } else {
    fc = 0.5 * cos(M_PI * r / Rc) + 0.5;
    dfc = -0.5 * M_PI / Rc * sin(M_PI * r / Rc);
  }
  *de_dr = dbase * fc + base * dfc;
  return base * fc;
}

/* Main entry.  Neighbours of LOCAL atoms come as CSR (nbr_off[nlocal+1], nbr[]).
 * species[ntotal] in 0..S-1 (ghosts keep their real species; they are "padding" only for the energy sum).
 * Outputs (Hartree, Hartree/Angstrom): energy, forces[ntotal*3], eatom[nlocal] (may be NULL),
 * virial9[9] row-major 3x3 (may be NULL), aev_out[nlocal*aev_len] (may be NULL),
 * daev_out[nlocal*aev_len] (may be NULL).  mode bit0: skip MLP/backward (AEV only).               */
int ani_oracle_compute(const omodel* m, int num_models_used, int radial_filter, int nlocal, int ntotal,
                       const int* species, const double* x, const int64_t* nbr_off, const int* nbr, double* energy,
                       double* forces, double* eatom, double* virial9, double* aev_out, double* daev_out, int aev_only,
                       int nthreads) {
  int nm = (num_models_used <= 0 || num_models_used > m->M) ? m->M : num_models_used;
  int L = m->aev_len;
  int maxnn = 0;
  for (int i = 0; i < nlocal; i++) {
    int c = (int)(nbr_off[i + 1] - nbr_off[i]);
    if (c > maxnn) maxnn = c;
  }
  int maxw = L;
  for (int s = 0; s < m->S; s++)
    for (int l = 0; l <= m->nlayers[s]; l++) if (m->dims[s][l] > maxw) maxw = m->dims[s][l];
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
  double* fthr = NULL;
  if (!aev_only) fthr = (double*)calloc((size_t)nthreads * ntotal * 3, sizeof(double));
  double* vthr = (double*)calloc((size_t)nthreads * 9, sizeof(double));
  double* ethr = (double*)calloc((size_t)nthreads, sizeof(double));
  int bad = 0;
#pragma omp parallel num_threads(nthreads)
  {
#ifdef _OPENMP
    int tid = omp_get_thread_num();
#else
    int tid = 0;
#endif
    onbr* nb = (onbr*)malloc(sizeof(onbr) * (maxnn + 1));
    double(*gvec)[3] = (double(*)[3])malloc(sizeof(double) * 3 * (maxnn + 1));
    double* G = (double*)malloc(sizeof(double) * L);
    double* dG = (double*)malloc(sizeof(double) * L);
    double* work = (double*)malloc(sizeof(double) * (size_t)(2 * MAXL + 2) * maxw);
    double* f = fthr ? fthr + (size_t)tid * ntotal * 3 : NULL;
    double* v = vthr + (size_t)tid * 9;
#pragma omp for schedule(dynamic, 16)
    for (int i = 0; i < nlocal; i++) {
      int si = species[i];
      if (si < 0 || si >= m->S) { bad = 1; continue; }
      if (m->use_rep && !aev_only) {      /* pair repulsion over the raw list entries (its own cutoff, no AEV filter) */
        for (int64_t p = nbr_off[i]; p < nbr_off[i + 1]; p++) {
          int j = nbr[p];
          if (species[j] < 0 || species[j] >= m->S) continue;
          double d[3] = {x[3 * j] - x[3 * i], x[3 * j + 1] - x[3 * i + 1], x[3 * j + 2] - x[3 * i + 2]};
          double r = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]), de;
          double e = rep_pair(m, si, species[j], r, &de);
          if (e == 0.0 && de == 0.0) continue;
          ethr[tid] += 0.5 * e;
          for (int c = 0; c < 3; c++) {
            double g = 0.5 * de * d[c] / r;     /* dE/dd_c */
            f[3 * i + c] += g;
            f[3 * j + c] -= g;
            for (int c2 = 0; c2 < 3; c2++) v[3 * c + c2] -= g * d[c2];
          }
        }
      }
      int nn = 0;
      for (int64_t p = nbr_off[i]; p < nbr_off[i + 1]; p++) {
        int j = nbr[p];
        onbr* q = &nb[nn];
        q->d[0] = x[3 * j] - x[3 * i]; q->d[1] = x[3 * j + 1] - x[3 * i + 1]; q->d[2] = x[3 * j + 2] - x[3 * i + 2];
        q->r = sqrt(q->d[0] * q->d[0] + q->d[1] * q->d[1] + q->d[2] * q->d[2]);
        q->j = j; q->s = species[j];
        if (q->s < 0 || q->s >= m->S) { bad = 1; continue; }
        /* pairs beyond the larger cutoff never contribute when the radial filter is on */
        if (radial_filter && q->r > m->Rcr && q->r > m->Rca) continue;
        nn++;
      }
      aev_row(m, radial_filter, nb, nn, G);
      if (aev_out) memcpy(aev_out + (size_t)i * L, G, sizeof(double) * L);
      if (aev_only) continue;
      double e = mlp_atom(m, nm, si, G, dG, work, maxw) + m->sae[si];
      if (daev_out) memcpy(daev_out + (size_t)i * L, dG, sizeof(double) * L);
      if (eatom) eatom[i] = e;
      ethr[tid] += e;
      aev_row_backward(m, radial_filter, nb, nn, dG, gvec);
      for (int a = 0; a < nn; a++) {
        int j = nb[a].j;
        for (int c = 0; c < 3; c++) {
          /* F = -dE/dx ; d = x_j - x_i  =>  F_i += g, F_j -= g */
          f[3 * i + c] += gvec[a][c];
          f[3 * j + c] -= gvec[a][c];
          for (int c2 = 0; c2 < 3; c2++) v[3 * c + c2] -= gvec[a][c] * nb[a].d[c2];
        }
      }
    }
    free(nb); free(gvec); free(G); free(dG); free(work);
  }
  if (!aev_only) {
    double e = 0.0;
    for (int t = 0; t < nthreads; t++) e += ethr[t];
    *energy = e;
    memset(forces, 0, sizeof(double) * (size_t)ntotal * 3);
    for (int t = 0; t < nthreads; t++) {
      const double* f = fthr + (size_t)t * ntotal * 3;
      for (size_t k = 0; k < (size_t)ntotal * 3; k++) forces[k] += f[k];
    }
    if (virial9) {
      double w[9] = {0};
      for (int t = 0; t < nthreads; t++) for (int k = 0; k < 9; k++) w[k] += vthr[t * 9 + k];
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) virial9[3 * a + b] = 0.5 * (w[3 * a + b] + w[3 * b + a]);
    }
  }
  free(fthr); free(vthr); free(ethr);
  return bad ? -1 : 0;
}
