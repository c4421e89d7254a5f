// jaco_render.cuh -- headless camera for the batched env: a CUDA ray caster over the model's analytic primitives.
//
// Replaces JacoMujocoEnvUtil._get_camera (env_script/env_mujoco_util.py:302-307: MjRenderContextOffscreen.render(640, 480, camera 0)
// + read_pixels(depth=True), both flipped top-down) for the fixed camera of jaco2_curtain_torque.xml:9,286.  The reference's OpenGL
// path (meshes, lights, materials) is unavailable on this box; this renderer draws what the simulation itself knows: the 33 collision
// primitives of the compiled model (plane / boxes / fitted capsules) plus two "bone" capsules for the upper-arm links that carry no
// collision geom (body origins 1-2 and 2-3), flat colour x head-light Lambert term, OpenGL-normalised depth like read_pixels.
//
// Two kernels:
//   jaco_render_prims_kernel  one thread per env: forward kinematics from the SoA state -> world pose of every primitive, then its
//                             camera-space record (16 floats: camera origin in the primitive's frame 3, R^T Rcam 9, size 3, type) in a
//                             scratch buffer, so that per pixel a primitive costs 6 FMA for the local ray direction + the hit test;
//   jaco_render_kernel        one 32x8 pixel tile of one env per block (a warp = one 32-pixel row segment): the env's records are staged
//                             in shared memory, culled against the tile's bounding cone (one thread per primitive), then every thread
//                             casts its ray against the surviving list.  Output: rgb u8 [env][H][W][3] (row 0 = top), depth f32 [env][H][W].
// Bound: instruction issue (about 8 surviving primitives x ~40 instructions per pixel); the HBM floor is 7 B per pixel.
// scripts/render_bench.py measures it.
#pragma once

#define JACO_RENDER_NBONE 2
#define JACO_RENDER_NPRIM (JM_NGEOM + JACO_RENDER_NBONE)
#define JACO_RENDER_TX 32
#define JACO_RENDER_TY 8

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
// camera-space record of a primitive (16 floats): ol[3] = camera origin in the primitive's frame, M[9] = R^T Rcam (row-major: a ray
// (x, y, -1) of the camera frame has the local direction M (x, y, -1)), size[3], type.  Everything per-pixel work needs; the world
// normal is not: the Lambert term n . d is invariant under the rotation into the primitive's frame.
JHD void render_to_camera(const JacoCamera& cam, const float* P, float* Q) {
  const float* R = P + 3;
  float r[3] = {cam.pos[0] - P[0], cam.pos[1] - P[1], cam.pos[2] - P[2]};
  m3tmulv(Q, R, r);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Q[3 + 3 * i + j] = R[i] * cam.R[j] + R[3 + i] * cam.R[3 + j] + R[6 + i] * cam.R[6 + j];
  Q[12] = P[12]; Q[13] = P[13]; Q[14] = P[14]; Q[15] = P[15];
}
// bounding-sphere radius of a primitive (plane: negative = unbounded)
JHD float render_prim_radius(const float* P) {
  const int type = (int)P[15];
  if (type == JACO_GEOM_PLANE) return -1.0f;
  if (type == JACO_GEOM_BOX) return sqrtf(P[12] * P[12] + P[13] * P[13] + P[14] * P[14]);
  return P[12] + P[13];
}
// local direction of the camera-frame ray (x, y, -1)
JHD void render_local_dir(const float* Q, float x, float y, float* dl) {
  const float* M = Q + 3;
  for (int k = 0; k < 3; k++) dl[k] = M[3 * k] * x + M[3 * k + 1] * y - M[3 * k + 2];
}
// ray ol + t dl (t = distance along the optical axis) against one primitive in its own frame; t in (tmin, tmax).  Returns true with
// the hit parameter and n . dl (n = outward unit normal at the hit) when it hits.
JHD bool render_hit(const float* Q, const float* dl, float tmin, float tmax, float* tout, float* ndl) {
  const float* ol = Q;
  const int type = (int)Q[15];
  float t = 0, nd = 0;
  if (type == JACO_GEOM_PLANE) {
    if (!(fabsf(dl[2]) > 1e-12f)) return false;
    t = -ol[2] / dl[2];
    if (!(t > tmin && t < tmax)) return false;
    nd = ol[2] > 0 ? dl[2] : -dl[2];
  } else if (type == JACO_GEOM_BOX) {
    float t0 = tmin, t1 = tmax; int ax = -1; float sg = 0;
    for (int k = 0; k < 3; k++) {
      const float h = Q[12 + k];
      if (fabsf(dl[k]) < 1e-12f) { if (fabsf(ol[k]) > h) return false; continue; }
      const float inv = 1.0f / dl[k];
      float ta = (-h - ol[k]) * inv, tb = (h - ol[k]) * inv, s_ = -1.0f;
      if (ta > tb) { float tt = ta; ta = tb; tb = tt; s_ = 1.0f; }
      if (ta > t0) { t0 = ta; ax = k; sg = s_; }
      if (tb < t1) t1 = tb;
      if (t0 > t1) return false;
    }
    if (ax < 0) return false;        // the ray starts inside the box (or before tmin inside it): not drawn
    t = t0; nd = sg * dl[ax];
  } else {                           // capsule: radius Q[12], half length Q[13] along local z
    const float rad = Q[12], h = Q[13], ir = 1.0f / rad;
    float best = tmax; bool hit = false;
    const float a = dl[0] * dl[0] + dl[1] * dl[1], b = ol[0] * dl[0] + ol[1] * dl[1], c = ol[0] * ol[0] + ol[1] * ol[1] - rad * rad;
    if (a > 1e-12f) {
      const float disc = b * b - a * c;
      if (disc >= 0) {
        const float tc = (-b - sqrtf(disc)) / a, z = ol[2] + tc * dl[2];
        if (tc > tmin && tc < best && fabsf(z) <= h) { best = tc; hit = true; nd = ((ol[0] + tc * dl[0]) * dl[0] + (ol[1] + tc * dl[1]) * dl[1]) * ir; }
      }
    }
    const float dd = a + dl[2] * dl[2];
    for (int e = -1; e <= 1; e += 2) {                  // the two end spheres
      const float oz = ol[2] - e * h, bb = b + oz * dl[2], cc = ol[0] * ol[0] + ol[1] * ol[1] + oz * oz - rad * rad, disc = bb * bb - dd * cc;
      if (disc < 0) continue;
      const float ts = (-bb - sqrtf(disc)) / dd, z = oz + ts * dl[2];
      if (ts > tmin && ts < best && z * e >= 0) { best = ts; hit = true; nd = ((ol[0] + ts * dl[0]) * dl[0] + (ol[1] + ts * dl[1]) * dl[1] + z * dl[2]) * ir; }
    }
    if (!hit) return false;
    t = best;
  }
  *tout = t; *ndl = nd;
  return true;
}
// pixel (row i from the top, column j) -> ray (x, y, -1) in the camera frame
JHD void render_pixel_xy(const JacoCamera& cam, int W, int H, float pj, float pi_, float* x, float* y) {
  const float ty = cam.tan_half_fovy, tx = ty * (float)W / (float)H;
  *x = (2 * pj / W - 1) * tx; *y = (1 - 2 * pi_ / H) * ty;
}
// nearest hit over a list of camera-space records -> colour and OpenGL-normalised depth
JHD void render_shade(const JacoCamera& cam, const float* recs, const int* list, int nlist, float x, float y, unsigned char* rgb, float* depth) {
  float best = cam.zfar, bnd = 0; int bid = -1;
  for (int k = 0; k < nlist; k++) {
    const int id = list ? list[k] : k;
    float t, nd, dl[3];
    render_local_dir(recs + 16 * id, x, y, dl);
    if (render_hit(recs + 16 * id, dl, cam.znear, cam.zfar, &t, &nd) && (t < best || (t == best && id < bid))) { best = t; bid = id; bnd = nd; }
  }
  if (bid < 0) { rgb[0] = rgb[1] = rgb[2] = 0; *depth = 1.0f; return; }
  float lam = -bnd / sqrtf(x * x + y * y + 1.0f);
  lam = cam.ambient + (1 - cam.ambient) * fmaxf(lam, 0.0f);
  for (int c = 0; c < 3; c++) rgb[c] = (unsigned char)(fminf(fmaxf(cam.rgb[bid][c] * lam, 0.0f), 1.0f) * 255.0f + 0.5f);
  *depth = (1.0f / cam.znear - 1.0f / best) / (1.0f / cam.znear - 1.0f / cam.zfar);
}

#ifdef __CUDACC__
__global__ void __launch_bounds__(128) jaco_render_prims_kernel(JacoCamera cam, const float* state, int N, int env_begin, int env_count, float* recs) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= env_count) return;
  JacoSim s;
  const int i = env_begin + e;
  for (int k = 0; k < JACO_NQ; k++) s.qpos[k] = state[(size_t)k * N + i];
  float P[JACO_RENDER_NPRIM * 16];
  render_prims(s, P);
  float* out = recs + (size_t)e * JACO_RENDER_NPRIM * 16;
  for (int g = 0; g < JACO_RENDER_NPRIM; g++) { float Q[16]; render_to_camera(cam, P + 16 * g, Q); for (int k = 0; k < 16; k++) out[16 * g + k] = Q[k]; }
}
// grid = (tiles_x, tiles_y, env_count); block = 32 x 8 threads: a warp is one 32-pixel row segment (128 B of depth, 96 B of rgb)
__global__ void __launch_bounds__(JACO_RENDER_TX * JACO_RENDER_TY) jaco_render_kernel(JacoCamera cam, const float* recs, int W, int H, unsigned char* rgb, float* depth) {
  __shared__ float sp[JACO_RENDER_NPRIM * 16];
  __shared__ int slist[JACO_RENDER_NPRIM];
  __shared__ unsigned int smask[2];
  const int tid = threadIdx.y * JACO_RENDER_TX + threadIdx.x, e = blockIdx.z;
  const float* P = recs + (size_t)e * JACO_RENDER_NPRIM * 16;
  for (int k = tid; k < JACO_RENDER_NPRIM * 16; k += JACO_RENDER_TX * JACO_RENDER_TY) sp[k] = P[k];
  if (tid < 2) smask[tid] = 0;
  __syncthreads();
  // tile culling: bounding sphere of primitive `tid` against the cone around the tile's centre ray that contains its four corners
  if (tid < JACO_RENDER_NPRIM) {
    const float j0 = blockIdx.x * JACO_RENDER_TX, i0 = blockIdx.y * JACO_RENDER_TY;
    const float j1 = fminf(j0 + JACO_RENDER_TX, (float)W), i1 = fminf(i0 + JACO_RENDER_TY, (float)H);
    float xc, yc, xk, yk;
    render_pixel_xy(cam, W, H, 0.5f * (j0 + j1), 0.5f * (i0 + i1), &xc, &yc);
    const float nc = sqrtf(xc * xc + yc * yc + 1.0f);
    float cosa = 1.0f;
    for (int c = 0; c < 4; c++) { render_pixel_xy(cam, W, H, (c & 1) ? j1 : j0, (c & 2) ? i1 : i0, &xk, &yk); cosa = fminf(cosa, (xc * xk + yc * yk + 1.0f) / (nc * sqrtf(xk * xk + yk * yk + 1.0f))); }
    cosa = fmaxf(cosa - 1e-4f, -1.0f);              // rounding slack
    const float* Q = sp + 16 * tid;
    const float rad = render_prim_radius(Q);
    bool keep = true;
    if (rad >= 0) {
      float dcl[3];
      render_local_dir(Q, xc, yc, dcl);             // tile-centre ray in the primitive's frame; the sphere centre is at -ol from the camera
      const float dist = v3norm(Q);
      if (dist > rad) {
        const float sinb = rad / dist, cosb = sqrtf(fmaxf(1 - sinb * sinb, 0.0f)), sina = sqrtf(fmaxf(1 - cosa * cosa, 0.0f));
        const float cos_ab = cosa * cosb - sina * sinb;           // cos(alpha + beta)
        const bool wide = (cosa * sinb + sina * cosb) < 0 || cos_ab < -0.99f;   // alpha + beta beyond pi: no culling
        keep = wide || -v3dot(Q, dcl) / (dist * nc) >= cos_ab - 1e-4f;
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
  const int j = blockIdx.x * JACO_RENDER_TX + threadIdx.x, i = blockIdx.y * JACO_RENDER_TY + threadIdx.y;
  if (j >= W || i >= H) return;
  float x, y;
  render_pixel_xy(cam, W, H, j + 0.5f, i + 0.5f, &x, &y);
  unsigned char c3[3]; float z;
  render_shade(cam, sp, slist, nl, x, y, c3, &z);
  const size_t pix = ((size_t)e * H + i) * W + j;
  rgb[3 * pix] = c3[0]; rgb[3 * pix + 1] = c3[1]; rgb[3 * pix + 2] = c3[2];
  depth[pix] = z;
}
#endif
