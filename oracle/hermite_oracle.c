/* TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called from the product path.
 *
 * CPU restatement ("port") of the reference's Hermite4CPU hot path in plain C.  Every function
 * cites the reference file:line it follows (paths relative to /root/reference/).  Expression
 * order, true divisions and float-typed quantities (Predictor.m, eta, e2 = s*s in float) are the
 * reference's, and the file is compiled like the reference (-O3 -fopenmp, x86-64 baseline, no FMA
 * contraction: oracle/Makefile) so that results are bit-identical to the reference's object code.
 *
 * Parity pin: the reference ships NO golden vectors or tests, so this
 * restatement is pinned against the reference itself, compiled unmodified into
 * oracle/_ref/libgravidy_ref.so (tests/test_oracle.py::test_port_matches_reference_*), and
 * against fixtures that library generated (tests/golden/, script tests/golden/make_golden.py).
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

/* reference: src/include/common.hpp:96-129 */
static const float GO_ETA_S = 0.01f;
static const double GO_D_TIME_MIN = 1.1920928955078125e-07; /* 2^-23 */
static const double GO_D_TIME_MAX = 0.125;                  /* 2^-3  */

/* reference: src/include/common.hpp:213-217 (mass is FLOAT) */
typedef struct { double r[3]; double v[3]; float m; } go_predictor;
/* reference: src/include/common.hpp:228-231 */
typedef struct { double a[3]; double a1[3]; } go_forces;
typedef struct { double x, y, z, w; } go_double4; /* common.hpp:67-70 */
typedef struct { double x, y, z; } go_double3;    /* common.hpp:72-75 */

typedef struct go_system
{
    uint32_t n;
    double e2;   /* NbodySystem.cpp:50: float*float product widened */
    float eta;   /* NbodySystem.hpp:67 */
    float total_mass, max_mass;
    go_double4 *h_r, *h_v;
    go_double3 *h_a2, *h_a3;
    go_forces *h_f, *h_old;
    go_predictor *h_p;
    double *h_t, *h_dt;
    uint32_t *h_move;
    double ekin, epot;
} go_system;

typedef struct go_run_stats
{
    uint64_t iterations, nsteps, interactions;
    double itime, atime, update_s, wall_s;
} go_run_stats;

/* reference: src/include/NbodySystem.cpp:36-69 (ctor), :161-190 (reader: m -> float),
 * :243-263 (copy_input_data), src/include/Hermite4.cpp:158-240 (alloc + init_data) */
go_system *go_create(uint32_t n, const double *m, const double *r3, const double *v3,
                     float softening, float eta)
{
    go_system *s = (go_system *)calloc(1, sizeof(go_system));
    s->n = n;
    s->e2 = softening * softening; /* float product, then widened */
    s->eta = eta;
    s->h_r = (go_double4 *)calloc(n, sizeof(go_double4));
    s->h_v = (go_double4 *)calloc(n, sizeof(go_double4));
    s->h_a2 = (go_double3 *)calloc(n, sizeof(go_double3));
    s->h_a3 = (go_double3 *)calloc(n, sizeof(go_double3));
    s->h_f = (go_forces *)calloc(n, sizeof(go_forces));
    s->h_old = (go_forces *)calloc(n, sizeof(go_forces));
    s->h_p = (go_predictor *)calloc(n, sizeof(go_predictor));
    s->h_t = (double *)calloc(n, sizeof(double));
    s->h_dt = (double *)calloc(n, sizeof(double));
    s->h_move = (uint32_t *)calloc(n, sizeof(uint32_t));
    s->total_mass = 0;
    s->max_mass = 0.0;
    for (uint32_t i = 0; i < n; i++)
    {
        float mf = (float)m[i]; /* NbodySystem.cpp:174 */
        if (mf > s->max_mass) s->max_mass = mf;
        s->total_mass += mf;    /* float accumulator, NbodySystem.cpp:190 */
        s->h_r[i].w = mf;
        s->h_r[i].x = r3[3*i]; s->h_r[i].y = r3[3*i+1]; s->h_r[i].z = r3[3*i+2];
        s->h_v[i].x = v3[3*i]; s->h_v[i].y = v3[3*i+1]; s->h_v[i].z = v3[3*i+2];
        /* Hermite4.cpp:198-210 */
        s->h_p[i].m = (float)s->h_r[i].w;
        s->h_p[i].r[0] = s->h_r[i].x; s->h_p[i].r[1] = s->h_r[i].y; s->h_p[i].r[2] = s->h_r[i].z;
        s->h_p[i].v[0] = s->h_v[i].x; s->h_p[i].v[1] = s->h_v[i].y; s->h_p[i].v[2] = s->h_v[i].z;
    }
    return s;
}

void go_destroy(go_system *s)
{
    if (!s) return;
    free(s->h_r); free(s->h_v); free(s->h_a2); free(s->h_a3); free(s->h_f); free(s->h_old);
    free(s->h_p); free(s->h_t); free(s->h_dt); free(s->h_move); free(s);
}

uint32_t go_n(go_system *s) { return s->n; }
double go_e2(go_system *s) { return s->e2; }
float go_eta(go_system *s) { return s->eta; }
float go_total_mass(go_system *s) { return s->total_mass; }
void go_set_threads(int nthreads) { omp_set_num_threads(nthreads); }
int go_max_threads(void) { return omp_get_max_threads(); }

/* reference: src/include/cpu/Hermite4CPU.cpp:55-82 */
static inline void go_force_calculation(const go_predictor *pi, const go_predictor *pj,
                                        go_forces *fi, double e2)
{
    double rx = pj->r[0] - pi->r[0];
    double ry = pj->r[1] - pi->r[1];
    double rz = pj->r[2] - pi->r[2];

    double vx = pj->v[0] - pi->v[0];
    double vy = pj->v[1] - pi->v[1];
    double vz = pj->v[2] - pi->v[2];

    double r2 = rx*rx + ry*ry + rz*rz + e2;
    double rinv = 1.0/sqrt(r2);
    double r2inv = rinv * rinv;
    double r3inv = r2inv * rinv;
    double r5inv = r2inv * r3inv;
    double mr3inv = r3inv * pj->m;
    double mr5inv = r5inv * pj->m;

    double rv = rx*vx + ry*vy + rz*vz;

    fi->a[0] += (rx * mr3inv);
    fi->a[1] += (ry * mr3inv);
    fi->a[2] += (rz * mr3inv);

    fi->a1[0] += (vx * mr3inv - (3 * rv) * rx * mr5inv);
    fi->a1[1] += (vy * mr3inv - (3 * rv) * ry * mr5inv);
    fi->a1[2] += (vz * mr3inv - (3 * rv) * rz * mr5inv);
}

/* exported single-pair entry for unit tests */
void go_pair(const double *pi7, const double *pj7, double e2, double *f6)
{
    go_predictor a, b; go_forces f;
    for (int k = 0; k < 3; k++) { a.r[k] = pi7[k]; a.v[k] = pi7[3+k]; b.r[k] = pj7[k]; b.v[k] = pj7[3+k]; }
    a.m = (float)pi7[6]; b.m = (float)pj7[6];
    for (int k = 0; k < 3; k++) { f.a[k] = f6[k]; f.a1[k] = f6[3+k]; }
    go_force_calculation(&a, &b, &f, e2);
    for (int k = 0; k < 3; k++) { f6[k] = f.a[k]; f6[3+k] = f.a1[k]; }
}

/* reference: src/include/cpu/Hermite4CPU.cpp:289-304 — accumulates INTO h_f (zeroed by init_data) */
void go_init_acc_jrk(go_system *s)
{
    #pragma omp parallel for
    for (uint32_t i = 0; i < s->n; i++)
        for (uint32_t j = 0; j < s->n; j++)
        {
            if (i == j) continue;
            go_force_calculation(&s->h_p[i], &s->h_p[j], &s->h_f[i], s->e2);
        }
}

/* reference: src/include/cpu/Hermite4CPU.cpp:309-336 (inner "parallel for" is inert: nesting off) */
void go_update_acc_jrk(go_system *s, uint32_t nact)
{
    #pragma omp parallel for
    for (uint32_t k = 0; k < nact; k++)
    {
        uint32_t i = s->h_move[k];
        memset(&s->h_f[i], 0, sizeof(go_forces));
        for (uint32_t j = 0; j < s->n; j++)
        {
            if (i == j) continue;
            go_force_calculation(&s->h_p[i], &s->h_p[j], &s->h_f[i], s->e2);
        }
    }
}

/* reference: src/include/cpu/Hermite4CPU.cpp:340-365 */
void go_predicted_pos_vel(go_system *s, double ITIME)
{
    for (uint32_t i = 0; i < s->n; i++)
    {
        double dt = ITIME - s->h_t[i];
        double dt2 = (dt * dt);
        double dt3 = (dt2 * dt);

        double dt2c = dt2/2.0;
        double dt3c = dt3/6.0;

        s->h_p[i].r[0] = (dt3c * s->h_f[i].a1[0]) + (dt2c * s->h_f[i].a[0]) + (dt * s->h_v[i].x) + s->h_r[i].x;
        s->h_p[i].r[1] = (dt3c * s->h_f[i].a1[1]) + (dt2c * s->h_f[i].a[1]) + (dt * s->h_v[i].y) + s->h_r[i].y;
        s->h_p[i].r[2] = (dt3c * s->h_f[i].a1[2]) + (dt2c * s->h_f[i].a[2]) + (dt * s->h_v[i].z) + s->h_r[i].z;

        s->h_p[i].v[0] = (dt2c * s->h_f[i].a1[0]) + (dt * s->h_f[i].a[0]) + s->h_v[i].x;
        s->h_p[i].v[1] = (dt2c * s->h_f[i].a1[1]) + (dt * s->h_f[i].a[1]) + s->h_v[i].y;
        s->h_p[i].v[2] = (dt2c * s->h_f[i].a1[2]) + (dt * s->h_f[i].a[2]) + s->h_v[i].z;

        s->h_p[i].m = (float)s->h_r[i].w;
    }
}

/* reference: src/include/utils/NbodyUtils.cpp:331-334 */
static inline double go_get_magnitude(double x, double y, double z)
{
    return sqrt(x*x + y*y + z*z);
}

/* reference: src/include/utils/NbodyUtils.cpp:337-365 */
double go_get_timestep_normal(go_system *s, uint32_t i, float ETA)
{
    double ax1_2 = s->h_a2[i].x + s->h_dt[i] * s->h_a3[i].x;
    double ay1_2 = s->h_a2[i].y + s->h_dt[i] * s->h_a3[i].y;
    double az1_2 = s->h_a2[i].z + s->h_dt[i] * s->h_a3[i].z;

    double abs_a1 = go_get_magnitude(s->h_f[i].a[0], s->h_f[i].a[1], s->h_f[i].a[2]);
    double abs_j1 = go_get_magnitude(s->h_f[i].a1[0], s->h_f[i].a1[1], s->h_f[i].a1[2]);
    double abs_j12 = abs_j1 * abs_j1;
    double abs_a1_3 = go_get_magnitude(s->h_a3[i].x, s->h_a3[i].y, s->h_a3[i].z);
    double abs_a1_2 = go_get_magnitude(ax1_2, ay1_2, az1_2);
    double abs_a1_22 = abs_a1_2 * abs_a1_2;

    double normal_dt = sqrt(ETA * ((abs_a1 * abs_a1_2 + abs_j12) / (abs_j1 * abs_a1_3 + abs_a1_22)));
    return normal_dt;
}

/* reference: src/include/utils/NbodyUtils.cpp:371-424 */
double go_normalize_dt(double new_dt, double old_dt, double t)
{
    if (new_dt <= old_dt/8)
    {
        new_dt = GO_D_TIME_MIN;
    }
    else if (old_dt/8 < new_dt && new_dt <= old_dt/4)
    {
        new_dt = old_dt / 8;
    }
    else if (old_dt/4 < new_dt && new_dt <= old_dt/2)
    {
        new_dt = old_dt / 4;
    }
    else if (old_dt/2 < new_dt && new_dt <= old_dt)
    {
        new_dt = old_dt / 2;
    }
    else if (old_dt < new_dt && new_dt <= old_dt * 2)
    {
        new_dt = old_dt;
    }
    else if (2 * old_dt < new_dt)
    {
        double val = t/(2 * old_dt);
        if (ceil(val) == val)
            new_dt = 2.0 * old_dt;
        else
            new_dt = old_dt;
    }
    else
    {
        new_dt = old_dt; /* reached for NaN */
    }

    if (new_dt < GO_D_TIME_MIN)
        new_dt = GO_D_TIME_MIN;
    else if (new_dt > GO_D_TIME_MAX)
        new_dt = GO_D_TIME_MAX;

    return new_dt;
}

/* reference: src/include/cpu/Hermite4CPU.cpp:370-411 */
void go_correction_pos_vel(go_system *s, double ITIME, uint32_t nact)
{
    for (uint32_t k = 0; k < nact; k++)
    {
        int i = s->h_move[k];

        double dt1 = s->h_dt[i];
        double dt2 = dt1 * dt1;
        double dt3 = dt2 * dt1;
        double dt4 = dt2 * dt2;
        double dt5 = dt4 * dt1;

        /* Acceleration 2nd derivative */
        s->h_a2[i].x = (-6 * (s->h_old[i].a[0] - s->h_f[i].a[0]) - dt1 * (4 * s->h_old[i].a1[0] + 2 * s->h_f[i].a1[0])) / dt2;
        s->h_a2[i].y = (-6 * (s->h_old[i].a[1] - s->h_f[i].a[1]) - dt1 * (4 * s->h_old[i].a1[1] + 2 * s->h_f[i].a1[1])) / dt2;
        s->h_a2[i].z = (-6 * (s->h_old[i].a[2] - s->h_f[i].a[2]) - dt1 * (4 * s->h_old[i].a1[2] + 2 * s->h_f[i].a1[2])) / dt2;

        /* Acceleration 3rd derivative */
        s->h_a3[i].x = (12 * (s->h_old[i].a[0] - s->h_f[i].a[0]) + 6 * dt1 * (s->h_old[i].a1[0] + s->h_f[i].a1[0])) / dt3;
        s->h_a3[i].y = (12 * (s->h_old[i].a[1] - s->h_f[i].a[1]) + 6 * dt1 * (s->h_old[i].a1[1] + s->h_f[i].a1[1])) / dt3;
        s->h_a3[i].z = (12 * (s->h_old[i].a[2] - s->h_f[i].a[2]) + 6 * dt1 * (s->h_old[i].a1[2] + s->h_f[i].a1[2])) / dt3;

        /* Correcting position */
        s->h_r[i].x = s->h_p[i].r[0] + (dt4/24)*s->h_a2[i].x + (dt5/120)*s->h_a3[i].x;
        s->h_r[i].y = s->h_p[i].r[1] + (dt4/24)*s->h_a2[i].y + (dt5/120)*s->h_a3[i].y;
        s->h_r[i].z = s->h_p[i].r[2] + (dt4/24)*s->h_a2[i].z + (dt5/120)*s->h_a3[i].z;

        /* Correcting velocity */
        s->h_v[i].x = s->h_p[i].v[0] + (dt3/6)*s->h_a2[i].x + (dt4/24)*s->h_a3[i].x;
        s->h_v[i].y = s->h_p[i].v[1] + (dt3/6)*s->h_a2[i].y + (dt4/24)*s->h_a3[i].y;
        s->h_v[i].z = s->h_p[i].v[2] + (dt3/6)*s->h_a2[i].z + (dt4/24)*s->h_a3[i].z;

        s->h_t[i] = ITIME;
        double normal_dt = go_get_timestep_normal(s, i, s->eta);
        normal_dt = go_normalize_dt(normal_dt, s->h_dt[i], s->h_t[i]);
        s->h_dt[i] = normal_dt;
    }
}

/* reference: src/include/Hermite4.cpp:63-84 (max_mass tracking at :72-73 kept) */
uint32_t go_find_particles_to_move(go_system *s, double ITIME)
{
    uint32_t j = 0;
    for (uint32_t i = 0; i < s->n; i++)
    {
        s->h_move[i] = (uint32_t)-1;
        if (s->h_r[i].w > s->max_mass) s->max_mass = s->h_r[i].w;
        double tmp_time = s->h_t[i] + s->h_dt[i];
        if (fabs(ITIME - tmp_time) < 2*DBL_EPSILON)
        {
            s->h_move[j] = i;
            j++;
        }
    }
    return j;
}

/* reference: src/include/Hermite4.cpp:90-102 */
double go_next_integration_time(go_system *s)
{
    double CTIME = s->h_t[0] + s->h_dt[0];
    for (uint32_t i = 0 + 1; i < s->n; i++)
    {
        double time = s->h_t[i] + s->h_dt[i];
        if (time < CTIME) CTIME = time;
    }
    return CTIME;
}

/* reference: src/include/Hermite4.cpp:108-140, called with ETA_S (Hermite4CPU_integration.cpp:54)
 * and CTIME preset to 1e10 (:43) */
double go_init_dt(go_system *s, double ITIME)
{
    double CTIME = 1.0e+10;
    float ETA = GO_ETA_S;
    for (uint32_t i = 0; i < s->n; i++)
    {
        go_forces f = s->h_f[i];
        double a2 = go_get_magnitude(f.a[0], f.a[1], f.a[2]);
        double j2 = go_get_magnitude(f.a1[0], f.a1[1], f.a1[2]);
        double tmp_dt = ETA * (a2/j2);
        int exp = (int)(ceil(log(tmp_dt)/log(2.0))-1);
        tmp_dt = pow(2, exp);
        if (tmp_dt < GO_D_TIME_MIN)
            tmp_dt = GO_D_TIME_MIN;
        else if (tmp_dt > GO_D_TIME_MAX)
            tmp_dt = GO_D_TIME_MAX;
        s->h_dt[i] = tmp_dt;
        s->h_t[i] = ITIME;
        if (tmp_dt < CTIME) CTIME = tmp_dt;
    }
    return CTIME;
}

/* reference: src/include/Hermite4.cpp:145-153 */
void go_save_old_acc_jrk(go_system *s, uint32_t nact)
{
    for (uint32_t k = 0; k < nact; k++)
    {
        uint32_t i = s->h_move[k];
        s->h_old[i] = s->h_f[i];
    }
}

/* reference: src/include/utils/NbodyUtils.cpp:427-529 (PN off).  The reference adds the per-i
 * partial sums with "omp atomic" in thread-arrival order; here they are added in ascending i, which
 * is the order the reference produces with one thread. */
double go_get_energy(go_system *s, double *ekin, double *epot)
{
    double *pe = (double *)malloc(sizeof(double) * s->n);
    double *ke = (double *)malloc(sizeof(double) * s->n);
    #pragma omp parallel for schedule(dynamic, 64)
    for (uint32_t i = 0; i < s->n; i++)
    {
        double epot_tmp = 0.0;
        for (uint32_t j = i+1; j < s->n; j++)
        {
            double rx = s->h_r[j].x - s->h_r[i].x;
            double ry = s->h_r[j].y - s->h_r[i].y;
            double rz = s->h_r[j].z - s->h_r[i].z;
            double r2 = rx*rx + ry*ry + rz*rz;
            epot_tmp -= (s->h_r[i].w * s->h_r[j].w) / sqrt(r2);
        }
        double vvx = s->h_v[i].x * s->h_v[i].x;
        double vvy = s->h_v[i].y * s->h_v[i].y;
        double vvz = s->h_v[i].z * s->h_v[i].z;
        double vv2 = vvx + vvy + vvz;
        ke[i] = 0.5 * s->h_r[i].w * vv2;
        pe[i] = epot_tmp;
    }
    s->ekin = 0.0; s->epot = 0.0;
    for (uint32_t i = 0; i < s->n; i++) { s->ekin += ke[i]; s->epot += pe[i]; }
    free(pe); free(ke);
    if (ekin) *ekin = s->ekin;
    if (epot) *epot = s->epot;
    return s->ekin + s->epot + 0.0;
}

void go_get_state(go_system *s, double *r4, double *v4, double *f6, double *old6, double *a2,
                  double *a3, double *t, double *dt, double *p7, uint32_t *move)
{
    for (uint32_t i = 0; i < s->n; i++)
    {
        if (r4) { r4[4*i] = s->h_r[i].x; r4[4*i+1] = s->h_r[i].y; r4[4*i+2] = s->h_r[i].z; r4[4*i+3] = s->h_r[i].w; }
        if (v4) { v4[4*i] = s->h_v[i].x; v4[4*i+1] = s->h_v[i].y; v4[4*i+2] = s->h_v[i].z; v4[4*i+3] = 0.0; }
        if (f6) for (int k = 0; k < 3; k++) { f6[6*i+k] = s->h_f[i].a[k]; f6[6*i+3+k] = s->h_f[i].a1[k]; }
        if (old6) for (int k = 0; k < 3; k++) { old6[6*i+k] = s->h_old[i].a[k]; old6[6*i+3+k] = s->h_old[i].a1[k]; }
        if (a2) { a2[3*i] = s->h_a2[i].x; a2[3*i+1] = s->h_a2[i].y; a2[3*i+2] = s->h_a2[i].z; }
        if (a3) { a3[3*i] = s->h_a3[i].x; a3[3*i+1] = s->h_a3[i].y; a3[3*i+2] = s->h_a3[i].z; }
        if (t) t[i] = s->h_t[i];
        if (dt) dt[i] = s->h_dt[i];
        if (p7) { for (int k = 0; k < 3; k++) { p7[7*i+k] = s->h_p[i].r[k]; p7[7*i+3+k] = s->h_p[i].v[k]; } p7[7*i+6] = s->h_p[i].m; }
        if (move) move[i] = s->h_move[i];
    }
}

void go_set_state(go_system *s, const double *r4, const double *v4, const double *f6,
                  const double *old6, const double *a2, const double *a3, const double *t,
                  const double *dt, const double *p7, const uint32_t *move)
{
    for (uint32_t i = 0; i < s->n; i++)
    {
        if (r4) { s->h_r[i].x = r4[4*i]; s->h_r[i].y = r4[4*i+1]; s->h_r[i].z = r4[4*i+2]; s->h_r[i].w = r4[4*i+3]; }
        if (v4) { s->h_v[i].x = v4[4*i]; s->h_v[i].y = v4[4*i+1]; s->h_v[i].z = v4[4*i+2]; }
        if (f6) for (int k = 0; k < 3; k++) { s->h_f[i].a[k] = f6[6*i+k]; s->h_f[i].a1[k] = f6[6*i+3+k]; }
        if (old6) for (int k = 0; k < 3; k++) { s->h_old[i].a[k] = old6[6*i+k]; s->h_old[i].a1[k] = old6[6*i+3+k]; }
        if (a2) { s->h_a2[i].x = a2[3*i]; s->h_a2[i].y = a2[3*i+1]; s->h_a2[i].z = a2[3*i+2]; }
        if (a3) { s->h_a3[i].x = a3[3*i]; s->h_a3[i].y = a3[3*i+1]; s->h_a3[i].z = a3[3*i+2]; }
        if (t) s->h_t[i] = t[i];
        if (dt) s->h_dt[i] = dt[i];
        if (p7) { for (int k = 0; k < 3; k++) { s->h_p[i].r[k] = p7[7*i+k]; s->h_p[i].v[k] = p7[7*i+3+k]; } s->h_p[i].m = (float)p7[7*i+6]; }
        if (move) s->h_move[i] = move[i];
    }
}

/* reference loop: src/include/cpu/Hermite4CPU_integration.cpp:81-124 without the file output.
 * 64-bit interaction counter (the reference's wraps at 2^32: SURVEY.md A.11). */
void go_run(go_system *s, double atime, double itime0, double t_end, uint64_t max_steps,
            go_run_stats *out, uint32_t *trace_nact, double *trace_itime, uint64_t trace_cap)
{
    double ATIME = atime;
    double ITIME = itime0;
    uint64_t it = 0, nsteps = 0, inter = 0;
    double t0 = omp_get_wtime(), upd = 0.0;
    while (ITIME < t_end)
    {
        if (max_steps && it >= max_steps) break;
        ITIME = ATIME;
        uint32_t nact = go_find_particles_to_move(s, ITIME);
        go_save_old_acc_jrk(s, nact);
        go_predicted_pos_vel(s, ITIME);
        double u0 = omp_get_wtime();
        go_update_acc_jrk(s, nact);
        upd += omp_get_wtime() - u0;
        go_correction_pos_vel(s, ITIME, nact);
        inter += (uint64_t)nact * (uint64_t)s->n;
        ATIME = go_next_integration_time(s);
        if (it < trace_cap)
        {
            if (trace_nact) trace_nact[it] = nact;
            if (trace_itime) trace_itime[it] = ITIME;
        }
        nsteps += nact;
        it++;
    }
    if (out)
    {
        out->iterations = it; out->nsteps = nsteps; out->interactions = inter;
        out->itime = ITIME; out->atime = ATIME; out->update_s = upd;
        out->wall_s = omp_get_wtime() - t0;
    }
}

/* Timed force sweep over an explicit active list (bench cpu_baseline "port" leg). */
double go_timed_update(go_system *s, const uint32_t *move, uint32_t nact)
{
    for (uint32_t k = 0; k < nact; k++) s->h_move[k] = move[k];
    double t0 = omp_get_wtime();
    go_update_acc_jrk(s, nact);
    return omp_get_wtime() - t0;
}
