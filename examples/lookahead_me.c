/* lookahead_me.c -- the call sequence of the headline path from plain C: a lookahead window of frame pairs
 * searched in ONE call of r1c_frame_me_window (full search +-range for every block of every size, SATD re-evaluation
 * and telescopic sub-pel refinement; what the lookahead of src/api/internal.rs:557-573 computes frame by frame).
 *
 *   gcc -std=c99 -Iinclude examples/lookahead_me.c -Lrav1e_b200 -lrav1e_cuda -Wl,-rpath,$PWD/rav1e_b200 -o lookahead_me
 *   ./lookahead_me [frames]          (needs a CUDA device; tests/test_cpp_mirror.py builds and runs it)
 *
 * The frames are synthetic: every frame is the previous one moved by (dx, dy) = (3, -2) pixels, so the search must
 * report the motion vector (row, col) = (-2 * 8, 3 * 8) in 1/8 pel for the blocks away from the frame border. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rav1e_cuda.h"

#define CHECK(call)                                                                    \
  do {                                                                                 \
    int rc_ = (call);                                                                  \
    if (rc_ != R1C_OK) {                                                               \
      fprintf(stderr, "%s failed: %d (%s)\n", #call, rc_, r1c_last_error(ctx));         \
      return 1;                                                                        \
    }                                                                                  \
  } while (0)

static unsigned lcg(unsigned *s) { return *s = *s * 1664525u + 1013904223u; }

int main(int argc, char **argv) {
  const int W = 320, H = 192, RANGE = 16, BD = 8;
  const int frames = argc > 1 ? atoi(argv[1]) : 4; /* pairs = frames - 1 */
  r1c_ctx *ctx = r1c_create(0);
  if (!ctx) {
    fprintf(stderr, "no CUDA device: librav1e_cuda has no CPU fallback\n");
    return 2;
  }
  /* Plane::new(W, H, 0, 0, 88, 88): the luma plane geometry of the encoder's frames */
  r1c_plane_cfg cfg;
  CHECK(r1c_plane_cfg_new(&cfg, W, H, 0, 0, 88, 88, 1));
  const size_t plane_bytes = (size_t)(cfg.stride * cfg.alloc_height);

  /* a smooth random field larger than the frame, from which every frame cuts a window moved by (3, -2) */
  const int FW = W + 64, FH = H + 64;
  unsigned char *field = malloc((size_t)FW * FH);
  unsigned seed = 7;
  for (int i = 0; i < FW * FH; i++) field[i] = (unsigned char)(lcg(&seed) >> 24);
  for (int pass = 0; pass < 3; pass++)
    for (int y = 1; y < FH - 1; y++)
      for (int x = 1; x < FW - 1; x++)
        field[y * FW + x] = (unsigned char)((field[y * FW + x] * 4 + field[y * FW + x - 1] + field[y * FW + x + 1] + field[(y - 1) * FW + x] +
                                             field[(y + 1) * FW + x] + 4) >> 3);

  r1c_plane **planes = calloc((size_t)frames, sizeof *planes);
  unsigned char *host = malloc(plane_bytes);
  for (int f = 0; f < frames; f++) {
    CHECK(r1c_plane_create(ctx, &cfg, 1, &planes[f]));
    memset(host, 0, plane_bytes);
    for (int y = 0; y < H; y++)
      memcpy(host + (size_t)(y + cfg.yorigin) * cfg.stride + cfg.xorigin, field + (size_t)(y + 32 - 2 * f) * FW + 32 + 3 * f, (size_t)W);
    CHECK(r1c_plane_upload(ctx, planes[f], host));
    CHECK(r1c_plane_pad(ctx, planes[f], W, H)); /* Plane::pad: the search may look past the frame edge */
  }

  r1c_frame_me_params p;
  memset(&p, 0, sizeof p);
  p.width = W, p.height = H, p.range_x = RANGE, p.range_y = RANGE, p.lambda = 40;
  p.bsize_mask = 0; /* every BlockSize up to 64x64 */
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++)
    if (r1c_block_width(b) <= 64 && r1c_block_height(b) <= 64) p.bsize_mask |= 1u << b;
  p.refine = 1, p.use_satd = 1;
  const long long n = r1c_frame_me_count(&p);

  /* frame f + 1 is searched in frame f: frame f + 1 shows the content of frame f moved by (-3, +2), i.e. the block at
   * (x, y) of frame f + 1 sits at (x + 3, y - 2) of frame f */
  const int pairs_n = frames - 1;
  r1c_frame_pair *pairs = calloc((size_t)pairs_n, sizeof *pairs);
  r1c_me_result *results = malloc((size_t)pairs_n * (size_t)n * sizeof *results);
  for (int i = 0; i < pairs_n; i++) {
    pairs[i].org = planes[i + 1];
    pairs[i].ref = planes[i];
    pairs[i].results = results + (size_t)i * (size_t)n;
  }
  CHECK(r1c_frame_me_window(ctx, pairs, pairs_n, &p, BD));

  int bad = 0, seen = 0;
  for (int i = 0; i < pairs_n; i++)
    for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
      int32_t cols, rows;
      const long long off = r1c_frame_me_offset(&p, b, &cols, &rows);
      if (off < 0) continue;
      const int bw = r1c_block_width(b), bh = r1c_block_height(b);
      for (int by = 0; by < rows; by++)
        for (int bx = 0; bx < cols; bx++) {
          if (bx * bw < 32 || by * bh < 32 || (bx + 1) * bw > W - 32 || (by + 1) * bh > H - 32) continue; /* border */
          const r1c_me_result *r = &results[(size_t)i * (size_t)n + (size_t)off + (size_t)by * cols + bx];
          seen++;
          bad += r->mv.row != -2 * 8 || r->mv.col != 3 * 8;
        }
    }
  printf("%d frame pairs, %lld blocks each, %lld kernel launches; %d interior blocks checked, %d off the true motion\n", pairs_n, n,
         (long long)r1c_launch_count(ctx), seen, bad);
  for (int f = 0; f < frames; f++) r1c_plane_destroy(planes[f]);
  r1c_destroy(ctx);
  free(planes), free(pairs), free(results), free(host), free(field);
  return bad * 50 > seen ? 1 : 0; /* flat spots of the field may tie: allow 2 % */
}
