/* synthgen.c — fast seeded ray caster for the large synthetic configs (C4: 2,000 keyframes, C5: 5 x 10^7 points).
 * TEST / BENCH INPUT GENERATION ONLY: no reference arithmetic lives here and nothing in the product links it.
 * numpy's scene.raycast does the same job for the small configs; 4 x 10^7 rays against a few hundred rectangles need C.
 * Deterministic for a given seed whatever the thread count: every scan draws from its own splitmix64 stream.
 * Build: gcc -O3 -march=native -fopenmp -shared -fPIC (tests/synthgen.py). */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct { double c[3], u[3], v[3], n[3], hu, hv; } rect_t;

static inline uint64_t splitmix64(uint64_t *s) {
  uint64_t z = (*s += 0x9e3779b97f4a7c15ull);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
static inline double uniform01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }
static inline double gauss(uint64_t *s) {
  double u1 = uniform01(s), u2 = uniform01(s);
  if (u1 < 1e-300) u1 = 1e-300;
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

/* rects: n_rects x 11 doubles (c3 u3 v3 hu hv, u and v unit and orthogonal); poses: n_scans x (R row-major 9 | p 3);
 * dirs: n_dirs x 3 unit body-frame directions, rotated per scan about body z by a random angle in [0, az_jitter).
 * out: n_scans x n_dirs x 3 floats (body-frame hits packed at the front of each scan's slab), counts: hits per scan. */
void synthgen_scans(const double *rects, int n_rects, const double *poses, int n_scans, const double *dirs, int n_dirs, double az_jitter,
                    double range_noise, double blind, double max_range, uint64_t seed, float *out, int32_t *counts) {
  rect_t *R = (rect_t *)malloc(sizeof(rect_t) * (size_t)n_rects);
  double *rad = (double *)malloc(sizeof(double) * (size_t)n_rects);
  for (int k = 0; k < n_rects; k++) {
    const double *r = rects + 11 * (size_t)k;
    for (int a = 0; a < 3; a++) { R[k].c[a] = r[a]; R[k].u[a] = r[3 + a]; R[k].v[a] = r[6 + a]; }
    R[k].hu = r[9]; R[k].hv = r[10];
    R[k].n[0] = R[k].u[1] * R[k].v[2] - R[k].u[2] * R[k].v[1];
    R[k].n[1] = R[k].u[2] * R[k].v[0] - R[k].u[0] * R[k].v[2];
    R[k].n[2] = R[k].u[0] * R[k].v[1] - R[k].u[1] * R[k].v[0];
    rad[k] = sqrt(R[k].hu * R[k].hu + R[k].hv * R[k].hv);
  }
#pragma omp parallel for schedule(dynamic, 4)
  for (int s = 0; s < n_scans; s++) {
    const double *P = poses + 12 * (size_t)s;
    const double *o = P + 9;
    uint64_t st = seed * 0x9e3779b97f4a7c15ull + (uint64_t)s * 0xd1b54a32d192ed03ull + 12345;
    const double az = az_jitter * uniform01(&st), ca = cos(az), sa = sin(az);
    /* rectangles within reach of this scan */
    int *near = (int *)malloc(sizeof(int) * (size_t)n_rects);
    int nn = 0;
    for (int k = 0; k < n_rects; k++) {
      const double dx = R[k].c[0] - o[0], dy = R[k].c[1] - o[1], dz = R[k].c[2] - o[2];
      if (sqrt(dx * dx + dy * dy + dz * dz) <= max_range + rad[k]) near[nn++] = k;
    }
    float *dst = out + 3 * (size_t)s * (size_t)n_dirs;
    int cnt = 0;
    for (int i = 0; i < n_dirs; i++) {
      const double *d0 = dirs + 3 * (size_t)i;
      const double db[3] = {ca * d0[0] - sa * d0[1], sa * d0[0] + ca * d0[1], d0[2]};
      const double dw[3] = {P[0] * db[0] + P[1] * db[1] + P[2] * db[2], P[3] * db[0] + P[4] * db[1] + P[5] * db[2], P[6] * db[0] + P[7] * db[1] + P[8] * db[2]};
      double best = INFINITY;
      for (int q = 0; q < nn; q++) {
        const rect_t *r = &R[near[q]];
        const double denom = dw[0] * r->n[0] + dw[1] * r->n[1] + dw[2] * r->n[2];
        if (fabs(denom) <= 1e-9) continue;
        const double t = ((r->c[0] - o[0]) * r->n[0] + (r->c[1] - o[1]) * r->n[1] + (r->c[2] - o[2]) * r->n[2]) / denom;
        if (!(t > 1e-6) || t >= best) continue;
        const double h[3] = {o[0] + t * dw[0] - r->c[0], o[1] + t * dw[1] - r->c[1], o[2] + t * dw[2] - r->c[2]};
        if (fabs(h[0] * r->u[0] + h[1] * r->u[1] + h[2] * r->u[2]) > r->hu) continue;
        if (fabs(h[0] * r->v[0] + h[1] * r->v[1] + h[2] * r->v[2]) > r->hv) continue;
        best = t;
      }
      const double g = gauss(&st);   /* drawn for every ray so that a hit's noise does not depend on its neighbours */
      if (!(best < INFINITY) || best <= blind || best > max_range) continue;
      const double t = best + range_noise * g;
      dst[3 * cnt] = (float)(db[0] * t); dst[3 * cnt + 1] = (float)(db[1] * t); dst[3 * cnt + 2] = (float)(db[2] * t);
      cnt++;
    }
    counts[s] = cnt;
    free(near);
  }
  free(R); free(rad);
}
