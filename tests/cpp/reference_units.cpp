// reference_units.cpp -- the reference's own unit tests for the hot path, restated against the C++ mirror
// (include/rav1e_cuda.hpp) of its interface:
//   host  (no GPU needed)  test_tx_log_scale and test_divu_pair (src/quantize.rs:160-203), get_mv_rate
//                          (src/me.rs:1005-1021), get_mv_range (:64-80), Plane::new geometry (src/frame/plane.rs:215-244),
//                          block / transform size tables, and the loud failure of Context without a device
//   gpu                    get_sad_same_u8 / _u16, get_satd_same_u8 / _u16 (src/dist.rs:341-465: the 22 known answers
//                          each), test_plane_pad (src/frame/plane.rs:707-752), roundtrips_u8 / _u16
//                          (src/transform/mod.rs:589-725: forward + inverse transform returns the source within the
//                          reference's per-combination tolerance), and a put_8tap / prep_8tap / mc_avg identity
// Built and run by tests/test_cpp_mirror.py:  g++ -std=c++17 -Iinclude reference_units.cpp -Lrav1e_b200 -lrav1e_cuda
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>

#include "rav1e_cuda.hpp"

static int g_checks = 0, g_failed = 0;
#define CHECK(cond)                                                      \
  do {                                                                   \
    g_checks++;                                                          \
    if (!(cond)) {                                                       \
      g_failed++;                                                        \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);      \
    }                                                                    \
  } while (0)
#define CHECK_EQ(a, b)                                                                                      \
  do {                                                                                                      \
    g_checks++;                                                                                             \
    const long long a_ = (long long)(a), b_ = (long long)(b);                                               \
    if (a_ != b_) {                                                                                         \
      g_failed++;                                                                                           \
      std::printf("FAILED %s:%d: %s == %s (%lld vs %lld)\n", __FILE__, __LINE__, #a, #b, a_, b_);           \
    }                                                                                                       \
  } while (0)

using namespace rav1e;

// ---- host ---------------------------------------------------------------------------------------------------
static void host_tests() {
  // get_mv_rate: 2 * bits of each component difference (halved first without high precision), src/me.rs:1005-1021
  const MotionVector z = {0, 0};
  CHECK_EQ(get_mv_rate(z, z, false), 0);
  CHECK_EQ(get_mv_rate(MotionVector{8, 0}, z, true), 2 * 4);   // |8| has 4 bits
  CHECK_EQ(get_mv_rate(MotionVector{8, 0}, z, false), 2 * 3);  // 8 >> 1 = 4: 3 bits
  CHECK_EQ(get_mv_rate(MotionVector{-8, 16}, z, false), 2 * 3 + 2 * 4);
  CHECK_EQ(get_mv_rate(MotionVector{1, 0}, z, false), 0);  // 1 >> 1 = 0
  CHECK_EQ(get_mv_rate(MotionVector{-1, 0}, z, false), 2);  // -1 >> 1 = -1 (arithmetic)
  // get_mv_range, src/me.rs:64-80: border = 128 + blk (in 1/8 px: 8 * ...), a block at the frame origin
  const MvRange r = get_mv_range(1920 / 4, 1080 / 4, 0, 0, 16, 16);
  CHECK_EQ(r.mvx_min, -(0 * 4 * 8) - 128 - 16 * 8);
  CHECK_EQ(r.mvy_min, -128 - 16 * 8);
  CHECK_EQ(r.mvx_max, (1920 / 4 - 0 - 16 / 4) * 4 * 8 + 128 + 16 * 8);
  CHECK_EQ(r.mvy_max, (1080 / 4 - 0 - 16 / 4) * 4 * 8 + 128 + 16 * 8);
  // Plane::new, src/frame/plane.rs:215-244: origin and stride aligned to 1 << (STRIDE_ALIGNMENT_LOG2 + 1 - size_of::<T>()) pixels
  for (int pxb = 1; pxb <= 2; pxb++) {
    const PlaneConfig c = plane_config_new(640, 480, 0, 0, 128 + 8, 128 + 8, pxb);
    const int64_t align = 1 << (5 + 1 - pxb);
    CHECK_EQ(c.xorigin, (136 + align - 1) / align * align);
    CHECK_EQ(c.yorigin, 136);
    CHECK_EQ(c.stride, (c.xorigin + 640 + 136 + align - 1) / align * align);
    CHECK_EQ(c.alloc_height, 136 + 480 + 136);
    CHECK_EQ(c.width, 640);
    CHECK_EQ(c.height, 480);
  }
  // BlockSize / TxSize tables (src/partition.rs, src/transform/mod.rs:225-247)
  CHECK_EQ(r1c_block_width(R1C_BLOCK_64X16), 64);
  CHECK_EQ(r1c_block_height(R1C_BLOCK_64X16), 16);
  CHECK_EQ(r1c_block_width(R1C_BLOCK_SIZES_ALL), -1);
  CHECK_EQ(r1c_tx_width(R1C_TX_16X64), 16);
  CHECK_EQ(r1c_tx_height(R1C_TX_16X64), 64);
  int valid = 0;
  for (int ts = 0; ts < 19; ts++)
    for (int tt = 0; tt < R1C_TX_TYPES; tt++) valid += r1c_fwd_txfm_valid(ts, tt);
  CHECK_EQ(valid, 193);  // the (size, type) pairs forward.rs:1702-1769 implements
  // test_tx_log_scale, src/quantize.rs:176-203
  static const struct {
    TxSize tx_size;
    int scale;
  } tx_sizes[] = {{R1C_TX_4X4, 0},   {R1C_TX_8X8, 0},   {R1C_TX_16X16, 0}, {R1C_TX_32X32, 1}, {R1C_TX_64X64, 2}, {R1C_TX_4X8, 0},  {R1C_TX_8X4, 0},
                  {R1C_TX_8X16, 0},  {R1C_TX_16X8, 0},  {R1C_TX_16X32, 1}, {R1C_TX_32X16, 1}, {R1C_TX_32X64, 2}, {R1C_TX_64X32, 2}, {R1C_TX_4X16, 0},
                  {R1C_TX_16X4, 0},  {R1C_TX_8X32, 0},  {R1C_TX_32X8, 0},  {R1C_TX_16X64, 1}, {R1C_TX_64X16, 1}};
  for (const auto &t : tx_sizes) CHECK_EQ(r1c_log_tx_scale(t.tx_size), t.scale);
  // test_divu_pair, src/quantize.rs:160-168: x / d == divu_pair(x, divu_gen(d)) with the (mul, add, shift) triples that
  // QuantizationContext::update leaves in the context the kernels divide with
  int bad = 0;
  for (int d = 1; d < 1024; d++) {
    r1c_quant_ctx qc;
    if (r1c_quant_ctx_update(&qc, R1C_TX_8X8, 0, (uint32_t)d, (uint32_t)d) != R1C_OK) {
      bad++;
      continue;
    }
    for (int x = -1000; x < 1000; x++) {
      const uint64_t y = (uint64_t)(x < 0 ? -x : x);
      const int q = (int)((((uint64_t)qc.dc_mul_add[0] * y + qc.dc_mul_add[1]) >> 32) >> qc.dc_mul_add[2]);
      bad += (x < 0 ? -q : q) != x / d;
      const int qa = (int)((((uint64_t)qc.ac_mul_add[0] * y + qc.ac_mul_add[1]) >> 32) >> qc.ac_mul_add[2]);
      bad += (x < 0 ? -qa : qa) != x / d;
    }
  }
  CHECK_EQ(bad, 0);
}

static void no_device_test() {
  // the product path fails loudly without a GPU: no CPU fallback behind this interface
  bool threw = false;
  try {
    Context ctx(0);
  } catch (const Error &e) {
    threw = e.code == R1C_ERR_CUDA;
  }
  CHECK(threw);
}

// ---- gpu ----------------------------------------------------------------------------------------------------
// setup_planes, src/dist.rs:341-375: two planes with different strides, a diagonal ramp each way
template <typename T>
static void setup_planes(Plane<T> &input_plane, Plane<T> &rec_plane) {
  const int xpad_off = (int)(input_plane.cfg.xorigin - input_plane.cfg.xpad) - 8;
  for (int64_t i = 0; i < input_plane.cfg.alloc_height; i++)
    for (int64_t j = 0; j < input_plane.cfg.stride; j++) input_plane.data[i * input_plane.cfg.stride + j] = (T)(((int)(j + i) - xpad_off) & 255);
  for (int64_t i = 0; i < rec_plane.cfg.alloc_height; i++)
    for (int64_t j = 0; j < rec_plane.cfg.stride; j++) rec_plane.data[i * rec_plane.cfg.stride + j] = (T)(((int)j - (int)i - xpad_off) & 255);
  input_plane.upload();
  rec_plane.upload();
}

struct Kat {
  BlockSize bsize;
  uint32_t sad, satd;
};
// the known answers of get_sad_same_inner / get_satd_same_inner, src/dist.rs:379-402, 437-460
static const Kat KATS[22] = {
    {R1C_BLOCK_4X4, 1912, 1408},        {R1C_BLOCK_4X8, 4296, 2016},        {R1C_BLOCK_8X4, 3496, 1816},
    {R1C_BLOCK_8X8, 7824, 3984},        {R1C_BLOCK_8X16, 16592, 5136},      {R1C_BLOCK_16X8, 14416, 4864},
    {R1C_BLOCK_16X16, 31136, 9984},     {R1C_BLOCK_16X32, 60064, 13824},    {R1C_BLOCK_32X16, 59552, 13760},
    {R1C_BLOCK_32X32, 120128, 27952},   {R1C_BLOCK_32X64, 186688, 37168},   {R1C_BLOCK_64X32, 250176, 45104},
    {R1C_BLOCK_64X64, 438912, 84176},   {R1C_BLOCK_64X128, 654272, 127920}, {R1C_BLOCK_128X64, 1016768, 173680},
    {R1C_BLOCK_128X128, 1689792, 321456}, {R1C_BLOCK_4X16, 8680, 3136},     {R1C_BLOCK_16X4, 6664, 2632},
    {R1C_BLOCK_8X32, 31056, 7056},      {R1C_BLOCK_32X8, 27600, 6624},      {R1C_BLOCK_16X64, 93344, 18432},
    {R1C_BLOCK_64X16, 116384, 21312}};

template <typename T>
static void get_sad_satd_same(const Context &ctx) {
  const int bit_depth = 8;
  Plane<T> input_plane(ctx, 640, 480, 0, 0, 128 + 8, 128 + 8);
  Plane<T> rec_plane(ctx, 640, 480, 0, 0, 2 * 128 + 8, 2 * 128 + 8);
  setup_planes(input_plane, rec_plane);
  for (const Kat &k : KATS) {
    // Area::StartingAt { x: 32, y: 40 }
    CHECK_EQ(get_sad(input_plane.region(32, 40), rec_plane.region(32, 40), k.bsize, bit_depth), k.sad);
    CHECK_EQ(get_satd(input_plane.region(32, 40), rec_plane.region(32, 40), k.bsize, bit_depth), k.satd);
  }
}

// test_plane_pad, src/frame/plane.rs:707-752
static void test_plane_pad(const Context &ctx) {
  PlaneConfig cfg = {};
  cfg.stride = 8, cfg.alloc_height = 9, cfg.width = 4, cfg.height = 4, cfg.xorigin = 2, cfg.yorigin = 3;
  Plane<uint8_t> plane(ctx, cfg, {0, 0, 0, 0, 0, 0, 0, 0,  //
                                  0, 0, 0, 0, 0, 0, 0, 0,  //
                                  0, 0, 0, 0, 0, 0, 0, 0,  //
                                  0, 0, 1, 2, 3, 4, 0, 0,  //
                                  0, 0, 8, 7, 6, 5, 0, 0,  //
                                  0, 0, 9, 8, 7, 6, 0, 0,  //
                                  0, 0, 2, 3, 4, 5, 0, 0,  //
                                  0, 0, 0, 0, 0, 0, 0, 0,  //
                                  0, 0, 0, 0, 0, 0, 0, 0});
  plane.pad(4, 4);
  const uint8_t want[72] = {1, 1, 1, 2, 3, 4, 4, 4,  //
                            1, 1, 1, 2, 3, 4, 4, 4,  //
                            1, 1, 1, 2, 3, 4, 4, 4,  //
                            1, 1, 1, 2, 3, 4, 4, 4,  //
                            8, 8, 8, 7, 6, 5, 5, 5,  //
                            9, 9, 9, 8, 7, 6, 6, 6,  //
                            2, 2, 2, 3, 4, 5, 5, 5,  //
                            2, 2, 2, 3, 4, 5, 5, 5,  //
                            2, 2, 2, 3, 4, 5, 5, 5};
  CHECK(std::memcmp(want, plane.data.data(), 72) == 0);
}

// test_roundtrip, src/transform/mod.rs:589-630
template <typename T>
static void test_roundtrip(const Context &ctx, std::mt19937 &rng, TxSize tx_size, TxType tx_type, int tolerance) {
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size), area = w * h;
  std::vector<T> src(area);
  PlaneConfig cfg = {};  // Plane::wrap(vec, stride = width)
  cfg.stride = w, cfg.alloc_height = h, cfg.width = w, cfg.height = h;
  std::vector<T> d(area);
  std::vector<int16_t> res(area);
  std::vector<int32_t> freq(area);
  for (int i = 0; i < area; i++) {
    src[i] = (T)(rng() & 255);
    d[i] = (T)(rng() & 255);
    res[i] = (int16_t)((int)src[i] - (int)d[i]);
  }
  Plane<T> dst(ctx, cfg, d);
  forward_transform(ctx, res, freq, tx_size, tx_type, 8);
  inverse_transform_add(freq, dst.region(0, 0), tx_size, tx_type, 8);
  int ne = 0;
  for (int i = 0; i < area; i++) ne += std::abs((int)src[i] - (int)dst.data[i]) > tolerance;
  if (ne) std::printf("roundtrip mismatch: tx_size %d tx_type %d tolerance %d: %d pixels\n", (int)tx_size, (int)tx_type, tolerance, ne);
  CHECK_EQ(ne, 0);
}

struct Combo {
  TxSize tx_size;
  TxType tx_type;
  int tolerance;
};
// `combinations` of roundtrips, src/transform/mod.rs:669-710
static const Combo COMBOS[] = {
    {R1C_TX_4X4, R1C_DCT_DCT, 0},     {R1C_TX_4X4, R1C_ADST_DCT, 0},    {R1C_TX_4X4, R1C_DCT_ADST, 0},   {R1C_TX_4X4, R1C_ADST_ADST, 0},
    {R1C_TX_4X4, R1C_IDTX, 0},        {R1C_TX_4X4, R1C_V_DCT, 0},       {R1C_TX_4X4, R1C_H_DCT, 0},      {R1C_TX_4X4, R1C_V_ADST, 0},
    {R1C_TX_4X4, R1C_H_ADST, 0},      {R1C_TX_8X8, R1C_DCT_DCT, 1},     {R1C_TX_8X8, R1C_ADST_DCT, 1},   {R1C_TX_8X8, R1C_DCT_ADST, 1},
    {R1C_TX_8X8, R1C_ADST_ADST, 1},   {R1C_TX_8X8, R1C_IDTX, 0},        {R1C_TX_8X8, R1C_V_DCT, 0},      {R1C_TX_8X8, R1C_H_DCT, 0},
    {R1C_TX_8X8, R1C_V_ADST, 0},      {R1C_TX_8X8, R1C_H_ADST, 1},      {R1C_TX_16X16, R1C_DCT_DCT, 1},  {R1C_TX_16X16, R1C_ADST_DCT, 1},
    {R1C_TX_16X16, R1C_DCT_ADST, 1},  {R1C_TX_16X16, R1C_ADST_ADST, 1}, {R1C_TX_16X16, R1C_IDTX, 0},     {R1C_TX_16X16, R1C_V_DCT, 1},
    {R1C_TX_16X16, R1C_H_DCT, 1},     {R1C_TX_32X32, R1C_DCT_DCT, 2},   {R1C_TX_32X32, R1C_IDTX, 0},     {R1C_TX_4X8, R1C_DCT_DCT, 1},
    {R1C_TX_8X4, R1C_DCT_DCT, 1},     {R1C_TX_4X16, R1C_DCT_DCT, 1},    {R1C_TX_16X4, R1C_DCT_DCT, 1},   {R1C_TX_8X16, R1C_DCT_DCT, 1},
    {R1C_TX_16X8, R1C_DCT_DCT, 1},    {R1C_TX_8X32, R1C_DCT_DCT, 2},    {R1C_TX_32X8, R1C_DCT_DCT, 2},   {R1C_TX_16X32, R1C_DCT_DCT, 2},
    {R1C_TX_32X16, R1C_DCT_DCT, 2}};

template <typename T>
static void roundtrips(const Context &ctx) {
  std::mt19937 rng(1234 + sizeof(T));
  for (const Combo &c : COMBOS)
    for (int rep = 0; rep < 4; rep++) test_roundtrip<T>(ctx, rng, c.tx_size, c.tx_type, c.tolerance);
}

// put_8tap with zero fractions copies the block (src/mc.rs:224-231); prep_8tap + mc_avg of one prediction with itself
// gives it back ((2 * (p << 4) + 16) >> 5 == p at 8 bits, :389-408)
template <typename T>
static void mc_identities(const Context &ctx, int bit_depth) {
  Plane<T> src(ctx, 128, 96, 0, 0, 16, 16), dst(ctx, 128, 96, 0, 0, 16, 16), avg(ctx, 128, 96, 0, 0, 16, 16);
  std::mt19937 rng(99);
  for (auto &v : src.data) v = (T)(rng() & ((1u << bit_depth) - 1));
  src.upload();
  put_8tap(dst.region(8, 12), src.region(40, 20), 32, 16, 0, 0, R1C_REGULAR, R1C_REGULAR, bit_depth);
  dst.download();
  const std::vector<int16_t> tmp = prep_8tap(src.region(40, 20), 32, 16, 0, 0, R1C_REGULAR, R1C_REGULAR, bit_depth);
  mc_avg(avg.region(8, 12), tmp, tmp, 32, 16, bit_depth);
  avg.download();
  int bad = 0;
  for (int y = 0; y < 16; y++)
    for (int x = 0; x < 32; x++) bad += dst.p(8 + x, 12 + y) != src.p(40 + x, 20 + y) || avg.p(8 + x, 12 + y) != src.p(40 + x, 20 + y);
  CHECK_EQ(bad, 0);
  CHECK_EQ(dst.p(7, 12), 128);  // nothing outside the block was written
  // a block that leaves the allocation is an error, not a silent clamp (the reference's slice bounds panic)
  bool threw = false;
  try {
    put_8tap(dst.region(8, 12), src.region(4000, 20), 32, 16, 0, 0, R1C_REGULAR, R1C_REGULAR, bit_depth);
  } catch (const Error &e) {
    threw = e.code == R1C_ERR_BOUNDS;
  }
  CHECK(threw);
}

int main(int argc, char **argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  host_tests();
  if (!gpu) {
    if (argc > 1 && std::strcmp(argv[1], "nodevice") == 0) no_device_test();
  } else {
    try {
      Context ctx(0);
      get_sad_satd_same<uint8_t>(ctx);
      get_sad_satd_same<uint16_t>(ctx);
      test_plane_pad(ctx);
      roundtrips<uint8_t>(ctx);
      roundtrips<uint16_t>(ctx);
      mc_identities<uint8_t>(ctx, 8);
      mc_identities<uint16_t>(ctx, 8);
      mc_identities<uint16_t>(ctx, 10);
    } catch (const Error &e) {
      std::printf("FAILED: rav1e::Error %d: %s\n", e.code, e.what());
      g_failed++;
    }
  }
  std::printf("%s: %d checks, %d failed\n", gpu ? "gpu" : "host", g_checks, g_failed);
  return g_failed ? 1 : 0;
}
