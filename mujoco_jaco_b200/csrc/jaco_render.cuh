// jaco_render.cuh -- headless camera for the batched env: a CUDA ray caster over the model's analytic primitives.
//
// Replaces JacoMujocoEnvUtil._get_camera (env_script/env_mujoco_util.py:302-307: MjRenderContextOffscreen.render(640, 480, camera 0)
// + read_pixels(depth=True), both flipped top-down) for the fixed camera of jaco2_curtain_torque.xml:9,286.  The reference's OpenGL
// path (meshes, lights, materials) is unavailable on this box; this renderer draws what the simulation itself knows: the 33 collision
// primitives of the compiled model (plane / boxes / fitted capsules) plus two "bone" capsules for the upper-arm links that carry no
// collision geom (body origins 1-2 and 2-3), flat colour x head-light Lambert term, OpenGL-normalised depth like read_pixels.
//
// Two kernels:
//   jaco_render_prims_kernel  one thread per env: forward kinematics from the SoA state -> world pose of every primitive
//                             (16 floats each: pos 3, R 9 row-major, size 3, type) in a scratch buffer;
//   jaco_render_kernel        one 16x16 pixel tile of one env per block: the env's primitives are staged in shared memory, culled
//                             against the tile's bounding cone (one thread per primitive), then every thread casts its ray against
//                             the surviving list.  Output: rgb u8 [env][H][W][3] (row 0 = top), depth f32 [env][H][W].
// Bound: HBM writes (7 B per pixel) once the tile culling has removed most of the 35 intersection tests; bench in scripts/render_bench.py.
#pragma once

#define JACO_RENDER_NBONE 2
#define JACO_RENDER_NPRIM (JM_NGEOM + JACO_RENDER_NBONE)
#define JACO_RENDER_TILE 16

struct JacoCamera {
  float pos[3], R[9];       // world position; columns of R = camera x (right), y (up), z (backwards: the camera looks along -z), row-major
  float tan_half_fovy;      // tan(fovy / 2)
  float znear, zfar;        // metres (MuJoCo: vis.map.znear / zfar x stat.extent)
  float rgb[JACO_RENDER_NPRIM][3];
  float ambient;            // colour = rgb * (ambient + (1 - ambient) * max(0, -n . d))
};

// frame with z = the given unit axis (the capsule is symmetric about z: the choice of x / y does not show)
JHD void render_frame_from_z(const float* z, float* R) {
  float a[3] = {1, 0, 0};
  if (fabsf(z[0]) > 0.9f) { a[0] = 0; a[1] = 1; }
  float x[3], y[3];
  v3cross(x, a, z); { float n = v3norm(x); x[0] /= n; x[1] /= n; x[2] /= n; }
  v3cross(y, z, x);
  for (int i = 0; i < 3; i++) { R[3 * i] = x[i]; R[3 * i + 1] = y[i]; R[3 * i + 2] = z[i]; }
}
// world primitives of one env from its qpos (fills prim[JACO_RENDER_NPRIM][16])
JHD void render_prims(JacoSim& s, float* prim) {
  const JacoModel& m = MDL;
  fk_chain(s);
  fk_free(s);
  for (int g = 0; g < JM_NGEOM; g++) {
    float* P = prim + 16 * g;
    geom_world(s, m.geoms[g], P, P + 3);
    P[12] = m.geoms[g].size[0]; P[13] = m.geoms[g].size[1]; P[14] = m.geoms[g].size[2]; P[15] = (float)m.geoms[g].type;
  }
  for (int b = 0; b < JACO_RENDER_NBONE; b++) {   // upper-arm links without collision geoms: capsule from body origin b+1 to b+2
    float* P = prim + 16 * (JM_NGEOM + b);
    float d[3]; v3sub(d, s.bx[b + 2], s.bx[b + 1]);
    float len = v3norm(d);
    if (len < 1e-6f) { d[0] = 0; d[1] = 0; d[2] = 1; len = 0; } else { d[0] /= len; d[1] /= len; d[2] /= len; }
    for (int i = 0; i < 3; i++) P[i] = 0.5f * (s.bx[b + 1][i] + s.bx[b + 2][i]);
    render_frame_from_z(d, P + 3);
    P[12] = 0.04f; P[13] = 0.5f * len; P[14] = 0; P[15] = (float)JACO_GEOM_CAPSULE;
  }
}
// bounding-sphere radius of a primitive (plane: negative = unbounded)
JHD float render_prim_radius(const float* P) {
  const int type = (int)P[15];
  if (type == JACO_GEOM_PLANE) return -1.0f;
  if (type == JACO_GEOM_BOX) return sqrtf(P[12] * P[12] + P[13] * P[13] + P[14] * P[14]);
  return P[12] + P[13];
}
// ray o + t d (d not normalised; t = distance along the optical axis) against one primitive; t in (tmin, tmax).  Returns true and
// the hit parameter / world normal when it hits.
JHD bool render_hit(const float* P, const float* o, const float* d, float tmin, float tmax, float* tout, float* nout) {
  const float* R = P + 3;
  float r[3] = {o[0] - P[0], o[1] - P[1], o[2] - P[2]}, ol[3], dl[3];
  m3tmulv(ol, R, r); m3tmulv(dl, R, d);
  const int type = (int)P[15];
  float t = 0, nl[3] = {0, 0, 0};
  if (type == JACO_GEOM_PLANE) {
    if (!(fabsf(dl[2]) > 1e-12f)) return false;
    t = -ol[2] / dl[2];
    if (!(t > tmin && t < tmax)) return false;
    nl[2] = ol[2] > 0 ? 1.0f : -1.0f;
  } else if (type == JACO_GEOM_BOX) {
    float t0 = tmin, t1 = tmax; int ax = -1; float sg = 0;
    for (int k = 0; k < 3; k++) {
      const float h = P[12 + k];
      if (fabsf(dl[k]) < 1e-12f) { if (fabsf(ol[k]) > h) return false; continue; }
      const float inv = 1.0f / dl[k];
      float ta = (-h - ol[k]) * inv, tb = (h - ol[k]) * inv, s_ = -1.0f;
      if (ta > tb) { float tt = ta; ta = tb; tb = tt; s_ = 1.0f; }
      if (ta > t0) { t0 = ta; ax = k; sg = s_; }
      if (tb < t1) t1 = tb;
      if (t0 > t1) return false;
    }
    if (ax < 0) return false;        // the ray starts inside the box (or before tmin inside it): not drawn
    t = t0; nl[ax] = sg;
  } else {                           // capsule: radius P[12], half length P[13] along local z
    const float rad = P[12], h = P[13];
    float best = tmax; bool hit = false;
    const float a = dl[0] * dl[0] + dl[1] * dl[1], b = ol[0] * dl[0] + ol[1] * dl[1], c = ol[0] * ol[0] + ol[1] * ol[1] - rad * rad;
    if (a > 1e-12f) {
      const float disc = b * b - a * c;
      if (disc >= 0) {
        const float tc = (-b - sqrtf(disc)) / a, z = ol[2] + tc * dl[2];
        if (tc > tmin && tc < best && fabsf(z) <= h) { best = tc; hit = true; nl[0] = (ol[0] + tc * dl[0]) / rad; nl[1] = (ol[1] + tc * dl[1]) / rad; nl[2] = 0; }
      }
    }
    const float dd = a + dl[2] * dl[2];
    for (int e = -1; e <= 1; e += 2) {                  // the two end spheres
      const float oz = ol[2] - e * h, bb = b + oz * dl[2], cc = ol[0] * ol[0] + ol[1] * ol[1] + oz * oz - rad * rad, disc = bb * bb - dd * cc;
      if (disc < 0) continue;
      const float ts = (-bb - sqrtf(disc)) / dd, z = oz + ts * dl[2];
      if (ts > tmin && ts < best && z * e >= 0) { best = ts; hit = true; nl[0] = (ol[0] + ts * dl[0]) / rad; nl[1] = (ol[1] + ts * dl[1]) / rad; nl[2] = z / rad; }
    }
    if (!hit) return false;
    t = best;
  }
  *tout = t;
  m3mulv(nout, R, nl);
  return true;
}
// pixel (row i from the top, column j) -> ray direction in the world, unnormalised, with unit component along the optical axis
JHD void render_ray(const JacoCamera& cam, int W, int H, float pj, float pi_, float* d) {
  const float ty = cam.tan_half_fovy, tx = ty * (float)W / (float)H;
  const float x = (2 * pj / W - 1) * tx, y = (1 - 2 * pi_ / H) * ty;
  for (int k = 0; k < 3; k++) d[k] = cam.R[3 * k] * x + cam.R[3 * k + 1] * y - cam.R[3 * k + 2];
}
// nearest hit over a primitive list -> colour and OpenGL-normalised depth
JHD void render_shade(const JacoCamera& cam, const float* prims, const int* list, int nlist, const float* d, unsigned char* rgb, float* depth) {
  float best = cam.zfar, bn[3] = {0, 0, 0}; int bid = -1;
  for (int k = 0; k < nlist; k++) {
    const int id = list ? list[k] : k;
    float t, n[3];
    if (render_hit(prims + 16 * id, cam.pos, d, cam.znear, cam.zfar, &t, n) && (t < best || (t == best && id < bid))) { best = t; bid = id; bn[0] = n[0]; bn[1] = n[1]; bn[2] = n[2]; }
  }
  if (bid < 0) { rgb[0] = rgb[1] = rgb[2] = 0; *depth = 1.0f; return; }
  const float dn = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  float lam = -(bn[0] * d[0] + bn[1] * d[1] + bn[2] * d[2]) / dn;
  lam = cam.ambient + (1 - cam.ambient) * fmaxf(lam, 0.0f);
  for (int c = 0; c < 3; c++) rgb[c] = (unsigned char)(fminf(fmaxf(cam.rgb[bid][c] * lam, 0.0f), 1.0f) * 255.0f + 0.5f);
  *depth = (1.0f / cam.znear - 1.0f / best) / (1.0f / cam.znear - 1.0f / cam.zfar);
}

#ifdef __CUDACC__
__global__ void __launch_bounds__(128) jaco_render_prims_kernel(const float* state, int N, int env_begin, int env_count, float* prims) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= env_count) return;
  JacoSim s;
  const int i = env_begin + e;
  for (int k = 0; k < JACO_NQ; k++) s.qpos[k] = state[(size_t)k * N + i];
  float P[JACO_RENDER_NPRIM * 16];
  render_prims(s, P);
  float* out = prims + (size_t)e * JACO_RENDER_NPRIM * 16;
  for (int k = 0; k < JACO_RENDER_NPRIM * 16; k++) out[k] = P[k];
}
// grid = (tiles_x, tiles_y, env_count); block = TILE x TILE threads
__global__ void __launch_bounds__(JACO_RENDER_TILE * JACO_RENDER_TILE) jaco_render_kernel(JacoCamera cam, const float* prims, int W, int H, unsigned char* rgb, float* depth) {
  __shared__ float sp[JACO_RENDER_NPRIM * 16];
  __shared__ int slist[JACO_RENDER_NPRIM];
  __shared__ unsigned int smask[2];
  const int tid = threadIdx.y * JACO_RENDER_TILE + threadIdx.x, e = blockIdx.z;
  const float* P = prims + (size_t)e * JACO_RENDER_NPRIM * 16;
  for (int k = tid; k < JACO_RENDER_NPRIM * 16; k += JACO_RENDER_TILE * JACO_RENDER_TILE) sp[k] = P[k];
  if (tid < 2) smask[tid] = 0;
  __syncthreads();
  // tile culling: bounding sphere of primitive `tid` against the cone around the tile's centre ray that contains its four corners
  if (tid < JACO_RENDER_NPRIM) {
    const float j0 = blockIdx.x * JACO_RENDER_TILE, i0 = blockIdx.y * JACO_RENDER_TILE;
    const float j1 = fminf(j0 + JACO_RENDER_TILE, (float)W), i1 = fminf(i0 + JACO_RENDER_TILE, (float)H);
    float dc[3], dk[3];
    render_ray(cam, W, H, 0.5f * (j0 + j1), 0.5f * (i0 + i1), dc);
    const float nc = v3norm(dc);
    float cosa = 1.0f;
    for (int c = 0; c < 4; c++) { render_ray(cam, W, H, (c & 1) ? j1 : j0, (c & 2) ? i1 : i0, dk); cosa = fminf(cosa, v3dot(dc, dk) / (nc * v3norm(dk))); }
    cosa = fmaxf(cosa - 1e-4f, -1.0f);              // rounding slack
    const float* Q = sp + 16 * tid;
    const float rad = render_prim_radius(Q);
    bool keep = true;
    if (rad >= 0) {
      float v[3] = {Q[0] - cam.pos[0], Q[1] - cam.pos[1], Q[2] - cam.pos[2]};
      const float dist = v3norm(v);
      if (dist > rad) {
        const float sinb = rad / dist, cosb = sqrtf(fmaxf(1 - sinb * sinb, 0.0f)), sina = sqrtf(fmaxf(1 - cosa * cosa, 0.0f));
        const float cos_ab = cosa * cosb - sina * sinb;           // cos(alpha + beta)
        const bool wide = (cosa * sinb + sina * cosb) < 0 || cos_ab < -0.99f;   // alpha + beta beyond pi: no culling
        keep = wide || v3dot(v, dc) / (dist * nc) >= cos_ab - 1e-4f;
      }
    }
    if (keep) atomicOr(&smask[tid >> 5], 1u << (tid & 31));
  }
  __syncthreads();
  if (tid == 0) {   // ordered list (ascending primitive id: deterministic)
    int n = 0;
    for (int k = 0; k < JACO_RENDER_NPRIM; k++) if (smask[k >> 5] >> (k & 31) & 1u) slist[n++] = k;
    smask[0] = (unsigned int)n;
  }
  __syncthreads();
  const int nl = (int)smask[0];
  const int j = blockIdx.x * JACO_RENDER_TILE + threadIdx.x, i = blockIdx.y * JACO_RENDER_TILE + threadIdx.y;
  if (j >= W || i >= H) return;
  float d[3];
  render_ray(cam, W, H, j + 0.5f, i + 0.5f, d);
  unsigned char c3[3]; float z;
  render_shade(cam, sp, slist, nl, d, c3, &z);
  const size_t pix = ((size_t)e * H + i) * W + j;
  rgb[3 * pix] = c3[0]; rgb[3 * pix + 1] = c3[1]; rgb[3 * pix + 2] = c3[2];
  depth[pix] = z;
}
#endif
