// rav1e_cuda.hpp -- C++17 host-side mirror of the reference's interface for the hot path, on top of the C ABI
// (rav1e_cuda.h).  Header only; link with librav1e_cuda.so.
//
// The reference is a Rust crate and the build image has no Rust toolchain, so the host side above the C ABI that a
// reader can compile and run here is this header: the SAME names, argument meaning and error behaviour as the
// reference's functions for the path, so that tests/cpp/ can restate the reference's own unit tests line by line
// (src/dist.rs:341-465 get_sad_same / get_satd_same, src/frame/plane.rs:707-752 test_plane_pad,
// src/transform/mod.rs:589-725 roundtrips).  The Rust crate a maintainer would add is bindings/rust/ +
// INTEGRATION.md; the Python mirror the GPU parity tests use is rav1e_b200/api.py.
//
//   reference                                                   here
//   Plane<T>::new(w, h, xdec, ydec, xpad, ypad)  plane.rs:215   rav1e::Plane<T>(ctx, w, h, xdec, ydec, xpad, ypad)
//   Plane { data, cfg } literal                  plane.rs:710   rav1e::Plane<T>(ctx, cfg, data)
//   plane.pad(w, h)                              plane.rs:269   plane.pad(w, h)
//   plane.downsample_from(&src)                  plane.rs:399   plane.downsample_from(src)
//   plane.region(Area::StartingAt { x, y })      tiling         plane.region(x, y)
//   get_sad(&org, &ref, bsize, bit_depth, cpu)   dist.rs:48     rav1e::get_sad(org_region, ref_region, bsize, bit_depth)
//   get_satd(...)                                dist.rs:121    rav1e::get_satd(...)
//   put_8tap / prep_8tap / mc_avg                mc.rs:410-706  rav1e::put_8tap / prep_8tap / mc_avg
//   forward_transform(res, coeffs, bd, ..)       mod.rs:556     rav1e::forward_transform(ctx, res, coeffs, tx_size, tx_type, bd)
//   inverse_transform_add(coeffs, &mut region..) mod.rs:568     rav1e::inverse_transform_add(coeffs, region, tx_size, tx_type, bd)
//   get_mv_rate / get_mv_range                   me.rs:64,1005  rav1e::get_mv_rate / get_mv_range
//   full_search                                  me.rs:943      rav1e::full_search
// Errors: the reference panics (assert!, unimplemented!, slice bounds); here every failing call throws rav1e::Error
// carrying the R1C_ERR_* code and the library's message.  There is no CPU fallback: without a CUDA device
// rav1e::Context's constructor throws.
#pragma once

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "rav1e_cuda.h"

namespace rav1e {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string &what) : std::runtime_error(what), code(c) {}
};

class Context {
 public:
  explicit Context(int device = 0) : h_(r1c_create(device)) {
    if (!h_) throw Error(R1C_ERR_CUDA, "no usable CUDA device: librav1e_cuda has no CPU fallback");
  }
  ~Context() { r1c_destroy(h_); }
  Context(const Context &) = delete;
  Context &operator=(const Context &) = delete;
  r1c_ctx *handle() const { return h_; }
  void check(int rc) const {
    if (rc != R1C_OK) throw Error(rc, r1c_last_error(h_));
  }
  void sync() const { check(r1c_sync(h_)); }

 private:
  r1c_ctx *h_;
};

// PlaneConfig, src/frame/plane.rs:22-47
using PlaneConfig = r1c_plane_cfg;
inline PlaneConfig plane_config_new(int64_t width, int64_t height, int64_t xdec, int64_t ydec, int64_t xpad, int64_t ypad,
                                    int px_bytes) {
  PlaneConfig c;
  if (r1c_plane_cfg_new(&c, width, height, xdec, ydec, xpad, ypad, px_bytes) != R1C_OK) throw Error(R1C_ERR_ARG, "bad plane geometry");
  return c;
}

template <typename T>
class Plane;

// PlaneRegion with Area::StartingAt { x, y }: a view of a plane from visible coordinate (x, y)
template <typename T>
struct PlaneRegion {
  Plane<T> *plane;
  int x, y;
};

// Plane<T>, src/frame/plane.rs:137-533: `data` is the whole padded allocation on the host (what the reference's tests
// fill and inspect), mirrored on the device; upload() / download() move it across.
template <typename T>
class Plane {
  static_assert(sizeof(T) == 1 || sizeof(T) == 2, "u8 or u16 pixels");

 public:
  std::vector<T> data;
  PlaneConfig cfg;

  Plane(const Context &ctx, int64_t width, int64_t height, int64_t xdec, int64_t ydec, int64_t xpad, int64_t ypad)
      : Plane(ctx, plane_config_new(width, height, xdec, ydec, xpad, ypad, (int)sizeof(T)), {}) {}
  // a plane given by its parts, like the `Plane { data, cfg }` literals of the reference's tests
  Plane(const Context &ctx, const PlaneConfig &c, std::vector<T> d) : data(std::move(d)), cfg(c), ctx_(&ctx), h_(nullptr) {
    if (data.empty()) data.assign((size_t)(cfg.stride * cfg.alloc_height), T(128));  // PlaneData::new, plane.rs:130-138
    if ((int64_t)data.size() != cfg.stride * cfg.alloc_height) throw Error(R1C_ERR_ARG, "plane data does not match its config");
    ctx.check(r1c_plane_create(ctx.handle(), &cfg, (int)sizeof(T), &h_));
    upload();
  }
  ~Plane() { r1c_plane_destroy(h_); }
  Plane(const Plane &) = delete;
  Plane &operator=(const Plane &) = delete;

  r1c_plane *handle() const { return h_; }
  const Context &context() const { return *ctx_; }
  void upload() { ctx_->check(r1c_plane_upload(ctx_->handle(), h_, data.data())); }
  void download() { ctx_->check(r1c_plane_download(ctx_->handle(), h_, data.data())); }
  // Plane::pad, src/frame/plane.rs:269-314 (on the device; `data` follows)
  void pad(int64_t w, int64_t h) {
    upload();
    ctx_->check(r1c_plane_pad(ctx_->handle(), h_, w, h));
    download();
  }
  // Plane::downsample_from, src/frame/plane.rs:399-423
  void downsample_from(Plane<T> &src) {
    src.upload();
    ctx_->check(r1c_plane_downsample(ctx_->handle(), h_, src.handle()));
    download();
  }
  PlaneRegion<T> region(int x, int y) { return PlaneRegion<T>{this, x, y}; }
  // pixel at visible coordinate (x, y)
  T &p(int x, int y) { return data[(size_t)((y + cfg.yorigin) * cfg.stride + x + cfg.xorigin)]; }

 private:
  const Context *ctx_;
  r1c_plane *h_;
};

using BlockSize = r1c_block_size;
using TxSize = r1c_tx_size;
using TxType = r1c_tx_type;
using FilterMode = r1c_filter_mode;
using MotionVector = r1c_mv;

// get_sad, src/dist.rs:48-111 (semantics of get_sad_ref :25-46)
template <typename T>
uint32_t get_sad(PlaneRegion<T> plane_org, PlaneRegion<T> plane_ref, BlockSize bsize, int bit_depth) {
  const Context &c = plane_org.plane->context();
  uint32_t out = 0;
  c.check(r1c_sad(c.handle(), plane_org.plane->handle(), plane_org.x, plane_org.y, plane_ref.plane->handle(), plane_ref.x,
                  plane_ref.y, bsize, bit_depth, &out));
  return out;
}
// get_satd, src/dist.rs:121-188 (get_satd_ref :197-328)
template <typename T>
uint32_t get_satd(PlaneRegion<T> plane_org, PlaneRegion<T> plane_ref, BlockSize bsize, int bit_depth) {
  const Context &c = plane_org.plane->context();
  uint32_t out = 0;
  c.check(r1c_satd(c.handle(), plane_org.plane->handle(), plane_org.x, plane_org.y, plane_ref.plane->handle(), plane_ref.x,
                   plane_ref.y, bsize, bit_depth, &out));
  return out;
}

// put_8tap, src/mc.rs:410-498: `src` region = the block's position in the reference plane (integer part of the MV
// already applied, like the PlaneSlice the reference passes), fractions in 1/16 px
template <typename T>
void put_8tap(PlaneRegion<T> dst, PlaneRegion<T> src, int width, int height, int col_frac, int row_frac, FilterMode mode_x,
              FilterMode mode_y, int bit_depth) {
  const Context &c = dst.plane->context();
  r1c_mc_job j = {dst.x, dst.y, src.x, src.y, (int16_t)width, (int16_t)height, (int8_t)col_frac, (int8_t)row_frac,
                  (int8_t)mode_x, (int8_t)mode_y, 0};
  c.check(r1c_put_8tap_batch(c.handle(), dst.plane->handle(), src.plane->handle(), &j, 1, bit_depth));
}
// prep_8tap, src/mc.rs:500-604: the i16 intermediate of a compound prediction, width * height values
template <typename T>
std::vector<int16_t> prep_8tap(PlaneRegion<T> src, int width, int height, int col_frac, int row_frac, FilterMode mode_x,
                               FilterMode mode_y, int bit_depth) {
  const Context &c = src.plane->context();
  std::vector<int16_t> tmp((size_t)width * height);
  r1c_mc_job j = {0, 0, src.x, src.y, (int16_t)width, (int16_t)height, (int8_t)col_frac, (int8_t)row_frac,
                  (int8_t)mode_x, (int8_t)mode_y, 0};
  c.check(r1c_prep_8tap_batch(c.handle(), tmp.data(), (int64_t)tmp.size(), src.plane->handle(), &j, 1, bit_depth));
  return tmp;
}
// mc_avg, src/mc.rs:606-706
template <typename T>
void mc_avg(PlaneRegion<T> dst, const std::vector<int16_t> &tmp1, const std::vector<int16_t> &tmp2, int width, int height,
            int bit_depth) {
  const Context &c = dst.plane->context();
  if (tmp1.size() != tmp2.size()) throw Error(R1C_ERR_ARG, "mc_avg: intermediates differ in size");
  r1c_avg_job j = {dst.x, dst.y, (int16_t)width, (int16_t)height, 0};
  c.check(r1c_mc_avg_batch(c.handle(), dst.plane->handle(), tmp1.data(), tmp2.data(), (int64_t)tmp1.size(), &j, 1, bit_depth));
}

// forward_transform, src/transform/mod.rs:556-566: residual W*H i16 (stride W) -> coeffs W*H i32
inline void forward_transform(const Context &c, const std::vector<int16_t> &residual, std::vector<int32_t> &coeffs, TxSize tx_size,
                              TxType tx_type, int bit_depth) {
  const size_t area = (size_t)r1c_tx_width(tx_size) * (size_t)r1c_tx_height(tx_size);
  if (residual.size() < area || coeffs.size() < area) throw Error(R1C_ERR_ARG, "forward_transform: buffers smaller than the block");
  c.check(r1c_forward_transform(c.handle(), residual.data(), coeffs.data(), tx_size, tx_type, bit_depth));
}
// inverse_transform_add, src/transform/mod.rs:568-580: adds the inverse transform of `coeffs` (the coded
// min(W,32) * min(H,32) coefficients, rows of min(W,32)) onto the W x H pixels of the region
template <typename T>
void inverse_transform_add(const std::vector<int32_t> &coeffs, PlaneRegion<T> output, TxSize tx_size, TxType tx_type, int bit_depth) {
  const Context &c = output.plane->context();
  r1c_itx_job j = {output.x, output.y};
  output.plane->upload();
  c.check(r1c_inv_txfm_add_batch(c.handle(), output.plane->handle(), &j, 1, tx_size, tx_type, bit_depth, coeffs.data()));
  output.plane->download();
}

// get_mv_rate, src/me.rs:1005-1021
inline uint32_t get_mv_rate(MotionVector a, MotionVector b, bool allow_high_precision_mv) {
  return r1c_mv_rate(a, b, allow_high_precision_mv ? 1 : 0);
}
// get_mv_range, src/me.rs:64-80 -> (mvx_min, mvx_max, mvy_min, mvy_max)
struct MvRange {
  int64_t mvx_min, mvx_max, mvy_min, mvy_max;
};
inline MvRange get_mv_range(int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h) {
  int64_t o[4];
  r1c_mv_range(w_in_b, h_in_b, bo_x, bo_y, blk_w, blk_h, o);
  return MvRange{o[0], o[1], o[2], o[3]};
}

// full_search, src/me.rs:943-990: raster scan of [x_lo, x_hi] x [y_lo, y_hi] in steps of `step`, strict `<`;
// returns the best MV and its cost (256 * sad + rate * lambda), u64::MAX when no candidate is legal
struct SearchResult {
  MotionVector mv;
  uint64_t cost;
};
template <typename T>
SearchResult full_search(Plane<T> &p_org, Plane<T> &p_ref, int po_x, int po_y, BlockSize bsize, int x_lo, int x_hi, int y_lo,
                         int y_hi, int step, uint32_t lambda, const MotionVector (&pmv)[2], bool allow_high_precision_mv,
                         int bit_depth) {
  const Context &c = p_org.context();
  r1c_fs_job j = {};
  j.po_x = po_x, j.po_y = po_y, j.bsize = bsize;
  j.x_lo = x_lo, j.x_hi = x_hi, j.y_lo = y_lo, j.y_hi = y_hi;
  j.step = step, j.lambda = lambda;
  j.pmv[0] = pmv[0], j.pmv[1] = pmv[1];
  j.allow_high_precision_mv = allow_high_precision_mv ? 1 : 0;
  r1c_me_result r = {};
  r.cost = ~0ull;
  c.check(r1c_full_search_batch(c.handle(), p_org.handle(), p_ref.handle(), &j, 1, bit_depth, &r));
  return SearchResult{r.mv, r.cost};
}

}  // namespace rav1e
