/*
 * jtsm_oracle.c -- CPU restatement of the reference's per-proposal hot path.
 *
 * THIS IS TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (jtsm_b200/) never imports, links or calls anything in oracle/.
 *
 * Plain C, single threaded unless the caller partitions the work (every entry
 * point takes a [roi_begin, roi_end) range so bench.py can fan out over host
 * threads).  Each function cites the reference file:line it follows; paths are
 * relative to the reference checkout (shenyunhang/JTSM).
 *
 * Parity pinning (see oracle/README.md, tests/test_oracle_*.py):
 *   - roi_align fwd:  golden vectors tests/layers/test_roi_align.py:26-39,
 *                     plus bit-equality with oracle/_ref (the reference's own
 *                     ROIAlign_cpu.cpp compiled in place) when available;
 *   - roi_pool, nms:  bit-equality with torchvision CPU ops (the third-party
 *                     dependency the reference calls; not pinned by any
 *                     reference test -> fixtures under tests/golden/);
 *   - pairwise_iou:   tests/structures/test_boxes.py:152-173;
 *   - matcher:        tests/modeling/test_matcher.py:19-23;
 *   - roi_loop_pool, moi_pool: parity unpinned by the reference (no test, no
 *                     runnable CPU path); restated line by line from the CUDA
 *                     kernels.
 *
 * Build: gcc -O2 -fPIC -shared -ffp-contract=off -o _build/libjtsm_oracle.so jtsm_oracle.c -lm
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------
 * RoIPool forward.  ROILoopPool_cpu.cpp:14-80 (in-tree copy of the
 * torchvision roi_pool algorithm the reference calls, poolers.py:183-186).
 * input [N,C,H,W], rois [R,5], output/argmax [R,C,PH,PW].
 * ------------------------------------------------------------------------ */
ORACLE_API void oracle_roi_pool_forward(const float* input, const float* rois, int channels,
                                        int height, int width, int pooled_h, int pooled_w,
                                        float spatial_scale, int roi_begin, int roi_end,
                                        float* output, int* argmax) {
  for (int n = roi_begin; n < roi_end; ++n) {
    const float* roi = rois + (size_t)n * 5;
    int b = (int)roi[0];
    int roi_start_w = (int)roundf(roi[1] * spatial_scale);
    int roi_start_h = (int)roundf(roi[2] * spatial_scale);
    int roi_end_w = (int)roundf(roi[3] * spatial_scale);
    int roi_end_h = (int)roundf(roi[4] * spatial_scale);
    int roi_width = imax(roi_end_w - roi_start_w + 1, 1);
    int roi_height = imax(roi_end_h - roi_start_h + 1, 1);
    float bin_h = (float)roi_height / (float)pooled_h;
    float bin_w = (float)roi_width / (float)pooled_w;
    for (int ph = 0; ph < pooled_h; ++ph) {
      for (int pw = 0; pw < pooled_w; ++pw) {
        int hstart = (int)floorf((float)ph * bin_h);
        int wstart = (int)floorf((float)pw * bin_w);
        int hend = (int)ceilf((float)(ph + 1) * bin_h);
        int wend = (int)ceilf((float)(pw + 1) * bin_w);
        hstart = imin(imax(hstart + roi_start_h, 0), height);
        hend = imin(imax(hend + roi_start_h, 0), height);
        wstart = imin(imax(wstart + roi_start_w, 0), width);
        wend = imin(imax(wend + roi_start_w, 0), width);
        int is_empty = (hend <= hstart) || (wend <= wstart);
        for (int c = 0; c < channels; ++c) {
          float maxval = is_empty ? 0.f : -FLT_MAX;
          int maxidx = -1;
          const float* in = input + ((size_t)b * channels + c) * height * width;
          for (int h = hstart; h < hend; ++h)
            for (int w = wstart; w < wend; ++w) {
              int idx = h * width + w;
              if (in[idx] > maxval) {
                maxval = in[idx];
                maxidx = idx;
              }
            }
          size_t o = (((size_t)n * channels + c) * pooled_h + ph) * pooled_w + pw;
          output[o] = maxval;
          argmax[o] = maxidx;
        }
      }
    }
  }
}

/* RoIPool / ROILoopPool / MOIPool backward: grad_input[argmax] += grad_output.
 * ROILoopPool_cpu.cpp:82-123; `n % num_rois` row->roi mapping for the stacked
 * ROILoopPool output follows ROILoopPool_cuda.cu:231.  grad_input must be
 * zero-initialised by the caller; rows [row_begin,row_end). */
ORACLE_API void oracle_max_pool_backward(const float* grad_output, const int* argmax,
                                         const float* rois, int num_rois, int channels, int height,
                                         int width, int pooled_h, int pooled_w, int row_begin,
                                         int row_end, float* grad_input) {
  const int pp = pooled_h * pooled_w;
  for (int n = row_begin; n < row_end; ++n) {
    int b = (int)rois[(size_t)(n % num_rois) * 5];
    for (int c = 0; c < channels; ++c) {
      float* gi = grad_input + ((size_t)b * channels + c) * height * width;
      const size_t base = ((size_t)n * channels + c) * pp;
      for (int i = 0; i < pp; ++i) {
        int a = argmax[base + i];
        if (a != -1) gi[a] += grad_output[base + i];
      }
    }
  }
}

/* ------------------------------------------------------------------------
 * RoIAlign.  detectron2/layers/csrc/ROIAlign/ROIAlign_cpu.cpp:21-114 (sample
 * pre-calc), :117-218 (forward), :220-395 (backward).  All arithmetic in
 * float, same operation order as the reference's T=float instantiation.
 * ------------------------------------------------------------------------ */
typedef struct {
  int pos1, pos2, pos3, pos4;
  float w1, w2, w3, w4;
} PreCalc;

/* ROIAlign_cpu.cpp:53-104 / 240-275: one bilinear sample -> 4 taps + weights.
 * Returns 0 if the sample is outside the map (all weights zero, taps -1). */
static int bilinear_taps(int height, int width, float y, float x, int* y_low, int* x_low,
                         int* y_high, int* x_high, float* w1, float* w2, float* w3, float* w4) {
  if (y < -1.0 || y > height || x < -1.0 || x > width) {
    *w1 = *w2 = *w3 = *w4 = 0.f;
    *x_low = *x_high = *y_low = *y_high = -1;
    return 0;
  }
  if (y <= 0) y = 0;
  if (x <= 0) x = 0;
  *y_low = (int)y;
  *x_low = (int)x;
  if (*y_low >= height - 1) {
    *y_high = *y_low = height - 1;
    y = (float)*y_low;
  } else {
    *y_high = *y_low + 1;
  }
  if (*x_low >= width - 1) {
    *x_high = *x_low = width - 1;
    x = (float)*x_low;
  } else {
    *x_high = *x_low + 1;
  }
  float ly = y - *y_low;
  float lx = x - *x_low;
  float hy = (float)(1. - ly), hx = (float)(1. - lx);
  *w1 = hy * hx;
  *w2 = hy * lx;
  *w3 = ly * hx;
  *w4 = ly * lx;
  return 1;
}

/* RoI geometry shared by forward and backward (ROIAlign_cpu.cpp:139-166). */
typedef struct {
  int batch;
  float start_h, start_w, bin_h, bin_w;
  int grid_h, grid_w;
} AlignGeom;

static AlignGeom align_geom(const float* roi, float spatial_scale, int pooled_h, int pooled_w,
                            int sampling_ratio, int aligned) {
  AlignGeom g;
  g.batch = (int)roi[0];
  float offset = aligned ? 0.5f : 0.0f;
  float roi_start_w = roi[1] * spatial_scale - offset;
  float roi_start_h = roi[2] * spatial_scale - offset;
  float roi_end_w = roi[3] * spatial_scale - offset;
  float roi_end_h = roi[4] * spatial_scale - offset;
  float roi_width = roi_end_w - roi_start_w;
  float roi_height = roi_end_h - roi_start_h;
  if (!aligned) {
    roi_width = fmaxf(roi_width, 1.f);
    roi_height = fmaxf(roi_height, 1.f);
  }
  g.start_h = roi_start_h;
  g.start_w = roi_start_w;
  g.bin_h = roi_height / (float)pooled_h;
  g.bin_w = roi_width / (float)pooled_w;
  g.grid_h = (sampling_ratio > 0) ? sampling_ratio : (int)ceilf(roi_height / pooled_h);
  g.grid_w = (sampling_ratio > 0) ? sampling_ratio : (int)ceilf(roi_width / pooled_w);
  return g;
}

ORACLE_API void oracle_roi_align_forward(const float* input, const float* rois, int channels,
                                         int height, int width, int pooled_h, int pooled_w,
                                         float spatial_scale, int sampling_ratio, int aligned,
                                         int roi_begin, int roi_end, float* output) {
  PreCalc* pre = NULL;
  size_t pre_cap = 0;
  for (int n = roi_begin; n < roi_end; ++n) {
    AlignGeom g = align_geom(rois + (size_t)n * 5, spatial_scale, pooled_h, pooled_w,
                             sampling_ratio, aligned);
    const int gh = g.grid_h > 0 ? g.grid_h : 0, gw = g.grid_w > 0 ? g.grid_w : 0;
    const float count = (float)imax(g.grid_h * g.grid_w, 1);
    size_t need = (size_t)gh * gw * pooled_h * pooled_w;
    if (need > pre_cap) {
      free(pre);
      pre = (PreCalc*)malloc((need ? need : 1) * sizeof(PreCalc));
      pre_cap = need;
    }
    size_t k = 0;
    for (int ph = 0; ph < pooled_h; ph++)
      for (int pw = 0; pw < pooled_w; pw++)
        for (int iy = 0; iy < gh; iy++) {
          const float yy = g.start_h + ph * g.bin_h + (float)(iy + .5f) * g.bin_h / (float)g.grid_h;
          for (int ix = 0; ix < gw; ix++) {
            const float xx = g.start_w + pw * g.bin_w + (float)(ix + .5f) * g.bin_w / (float)g.grid_w;
            int yl, xl, yh, xh;
            PreCalc pc;
            if (!bilinear_taps(height, width, yy, xx, &yl, &xl, &yh, &xh, &pc.w1, &pc.w2, &pc.w3, &pc.w4)) {
              pc.pos1 = pc.pos2 = pc.pos3 = pc.pos4 = 0;
            } else {
              pc.pos1 = yl * width + xl;
              pc.pos2 = yl * width + xh;
              pc.pos3 = yh * width + xl;
              pc.pos4 = yh * width + xh;
            }
            pre[k++] = pc;
          }
        }
    for (int c = 0; c < channels; c++) {
      const float* in = input + ((size_t)g.batch * channels + c) * height * width;
      float* out = output + ((size_t)n * channels + c) * pooled_h * pooled_w;
      k = 0;
      for (int b = 0; b < pooled_h * pooled_w; b++) {
        float v = 0.f;
        for (int s = 0; s < gh * gw; s++) {
          PreCalc pc = pre[k++];
          v += pc.w1 * in[pc.pos1] + pc.w2 * in[pc.pos2] + pc.w3 * in[pc.pos3] + pc.w4 * in[pc.pos4];
        }
        out[b] = v / count;
      }
    }
  }
  free(pre);
}

/* ROIAlign_cpu.cpp:287-395.  grad_input zero-initialised by the caller. */
ORACLE_API void oracle_roi_align_backward(const float* grad_output, const float* rois, int channels,
                                          int height, int width, int pooled_h, int pooled_w,
                                          float spatial_scale, int sampling_ratio, int aligned,
                                          int roi_begin, int roi_end, float* grad_input) {
  for (int n = roi_begin; n < roi_end; ++n) {
    AlignGeom g = align_geom(rois + (size_t)n * 5, spatial_scale, pooled_h, pooled_w,
                             sampling_ratio, aligned);
    const float count = (float)(g.grid_h * g.grid_w);
    for (int c = 0; c < channels; c++) {
      float* gi = grad_input + ((size_t)g.batch * channels + c) * height * width;
      const float* go = grad_output + ((size_t)n * channels + c) * pooled_h * pooled_w;
      for (int ph = 0; ph < pooled_h; ph++)
        for (int pw = 0; pw < pooled_w; pw++) {
          const float gb = go[ph * pooled_w + pw];
          for (int iy = 0; iy < g.grid_h; iy++) {
            const float y = g.start_h + ph * g.bin_h + (float)(iy + .5f) * g.bin_h / (float)g.grid_h;
            for (int ix = 0; ix < g.grid_w; ix++) {
              const float x = g.start_w + pw * g.bin_w + (float)(ix + .5f) * g.bin_w / (float)g.grid_w;
              float w1, w2, w3, w4;
              int xl, xh, yl, yh;
              bilinear_taps(height, width, y, x, &yl, &xl, &yh, &xh, &w1, &w2, &w3, &w4);
              float g1 = gb * w1 / count, g2 = gb * w2 / count, g3 = gb * w3 / count, g4 = gb * w4 / count;
              if (xl >= 0 && xh >= 0 && yl >= 0 && yh >= 0) {
                gi[yl * width + xl] += g1;
                gi[yl * width + xh] += g2;
                gi[yh * width + xl] += g3;
                gi[yh * width + xh] += g4;
              }
            }
          }
        }
    }
  }
}

/* ------------------------------------------------------------------------
 * ROILoopPool forward.  projects/WSL/wsl/layers/csrc/ROILoopPool/
 * ROILoopPool_cuda.cu:24-203 restated as host loops (the reference has no
 * runnable CPU path: ROILoopPool.h:62-64 raises before reaching it).
 * output/argmax are [3*R, C, PH, PW]: rows [0,R) = RoI, [R,2R) = frame,
 * [2R,3R) = context.  context_ratio is 1.8 at the only call site (:309).
 * maxval starts at 0 for all three pools (:109-110,178).
 * ------------------------------------------------------------------------ */
static float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

ORACLE_API void oracle_roi_loop_pool_forward(const float* input, const float* rois, int num_rois,
                                             int channels, int height, int width, int pooled_h,
                                             int pooled_w, float spatial_scale, float context_ratio,
                                             int roi_begin, int roi_end, float* output,
                                             int* argmax) {
  const size_t set = (size_t)num_rois * channels * pooled_h * pooled_w;
  for (int n = roi_begin; n < roi_end; ++n) {
    const float* roi = rois + (size_t)n * 5;
    int b = (int)roi[0];
    float x1 = roi[1], y1 = roi[2], x2 = roi[3], y2 = roi[4];
    float rois_w = x2 - x1, rois_h = y2 - y1;
    float rois_inner_w = rois_w / context_ratio, rois_inner_h = rois_h / context_ratio;
    float rois_outer_w = rois_w * context_ratio, rois_outer_h = rois_h * context_ratio;
    float inner_residual_w = rois_w - rois_inner_w, inner_residual_h = rois_h - rois_inner_h;
    float outer_residual_w = rois_outer_w - rois_w, outer_residual_h = rois_outer_h - rois_h;
    float x1_inner = x1 + inner_residual_w / 2, y1_inner = y1 + inner_residual_h / 2;
    float x2_inner = x2 - inner_residual_w / 2, y2_inner = y2 - inner_residual_h / 2;
    float x1_outer = x1 - outer_residual_w / 2, y1_outer = y1 - outer_residual_h / 2;
    float x2_outer = x2 + outer_residual_w / 2, y2_outer = y2 + outer_residual_h / 2;
    /* T(1.0 * width / spatial_scale): double division, then narrowed to T */
    float lim_w = (float)(1.0 * width / spatial_scale), lim_h = (float)(1.0 * height / spatial_scale);
    x1_inner = clampf(x1_inner, 0.f, lim_w); y1_inner = clampf(y1_inner, 0.f, lim_h);
    x2_inner = clampf(x2_inner, 0.f, lim_w); y2_inner = clampf(y2_inner, 0.f, lim_h);
    x1_outer = clampf(x1_outer, 0.f, lim_w); y1_outer = clampf(y1_outer, 0.f, lim_h);
    x2_outer = clampf(x2_outer, 0.f, lim_w); y2_outer = clampf(y2_outer, 0.f, lim_h);

    for (int pass = 0; pass < 2; ++pass) {
      int roi_start_w, roi_start_h, roi_end_w, roi_end_h, sw_in, sh_in, ew_in, eh_in;
      if (pass == 0) { /* RoI + frame (:75-142) */
        roi_start_w = (int)roundf(roi[1] * spatial_scale); roi_start_h = (int)roundf(roi[2] * spatial_scale);
        roi_end_w = (int)roundf(roi[3] * spatial_scale);   roi_end_h = (int)roundf(roi[4] * spatial_scale);
        sw_in = (int)roundf(x1_inner * spatial_scale); sh_in = (int)roundf(y1_inner * spatial_scale);
        ew_in = (int)roundf(x2_inner * spatial_scale); eh_in = (int)roundf(y2_inner * spatial_scale);
      } else { /* context (:144-202) */
        roi_start_w = (int)roundf(x1_outer * spatial_scale); roi_start_h = (int)roundf(y1_outer * spatial_scale);
        roi_end_w = (int)roundf(x2_outer * spatial_scale);   roi_end_h = (int)roundf(y2_outer * spatial_scale);
        sw_in = (int)roundf(roi[1] * spatial_scale); sh_in = (int)roundf(roi[2] * spatial_scale);
        ew_in = (int)roundf(roi[3] * spatial_scale); eh_in = (int)roundf(roi[4] * spatial_scale);
      }
      int roi_width = imax(roi_end_w - roi_start_w + 1, 1);
      int roi_height = imax(roi_end_h - roi_start_h + 1, 1);
      float bin_h = (float)roi_height / (float)pooled_h;
      float bin_w = (float)roi_width / (float)pooled_w;
      for (int ph = 0; ph < pooled_h; ++ph)
        for (int pw = 0; pw < pooled_w; ++pw) {
          int hstart = (int)floorf((float)ph * bin_h), wstart = (int)floorf((float)pw * bin_w);
          int hend = (int)ceilf((float)(ph + 1) * bin_h), wend = (int)ceilf((float)(pw + 1) * bin_w);
          hstart = imin(imax(hstart + roi_start_h, 0), height);
          hend = imin(imax(hend + roi_start_h, 0), height);
          wstart = imin(imax(wstart + roi_start_w, 0), width);
          wend = imin(imax(wend + roi_start_w, 0), width);
          for (int c = 0; c < channels; ++c) {
            const float* in = input + ((size_t)b * channels + c) * height * width;
            float maxval = 0.f, maxval_f = 0.f;
            int maxidx = -1, maxidx_f = -1;
            for (int h = hstart; h < hend; ++h)
              for (int w = wstart; w < wend; ++w) {
                int idx = h * width + w;
                if (pass == 0 && in[idx] > maxval) { maxval = in[idx]; maxidx = idx; }
                int inside_h = h > sh_in && h < eh_in;
                int inside_w = w > sw_in && w < ew_in;
                if (inside_h && inside_w) continue;
                if (in[idx] > maxval_f) { maxval_f = in[idx]; maxidx_f = idx; }
              }
            size_t o = (((size_t)n * channels + c) * pooled_h + ph) * pooled_w + pw;
            if (pass == 0) {
              output[o] = maxval;          argmax[o] = maxidx;
              output[o + set] = maxval_f;  argmax[o + set] = maxidx_f;
            } else {
              output[o + 2 * set] = maxval_f; argmax[o + 2 * set] = maxidx_f;
            }
          }
        }
    }
  }
}

/* ------------------------------------------------------------------------
 * MOIPool forward.  projects/WSL/wsl/layers/csrc/MOIPool/MOIPool_cuda.cu:
 * MoIForward :139-215 (mask), MoIPoolForward2 :218-294 (masked max).
 * oh_labels [R,S] int32, superpixels [N,Hs,Ws] int32.  `mois` is scratch
 * [H*W] ints per call.
 * ------------------------------------------------------------------------ */
ORACLE_API void oracle_moi_pool_forward(const float* input, const float* rois, const int* oh_labels,
                                        int len_labels, const int* superpixels, int height_sp,
                                        int width_sp, int channels, int height, int width,
                                        int pooled_h, int pooled_w, float spatial_scale,
                                        int roi_begin, int roi_end, float* output, int* argmax) {
  int* mois = (int*)malloc((size_t)height * width * sizeof(int));
  for (int n = roi_begin; n < roi_end; ++n) {
    const float* roi = rois + (size_t)n * 5;
    int b = (int)roi[0];
    int roi_start_w = (int)roundf(roi[1] * spatial_scale);
    int roi_start_h = (int)roundf(roi[2] * spatial_scale);
    int roi_end_w = (int)roundf(roi[3] * spatial_scale);
    int roi_end_h = (int)roundf(roi[4] * spatial_scale);
    const int* spp = superpixels + (size_t)b * height_sp * width_sp;
    const int* ohl = oh_labels + (size_t)n * len_labels;
    /* T spatial_scale_sp = 1.0 * height / height_sp  (:175) */
    float ss = (float)(1.0 * height / height_sp);
    for (int h = 0; h < height; ++h)
      for (int w = 0; w < width; ++w) {
        int v = 0;
        if (w >= roi_start_w && w <= roi_end_w && h >= roi_start_h && h <= roi_end_h) {
          int hhstart = (int)floorf((float)h / ss), wwstart = (int)floorf((float)w / ss);
          int hhend = (int)ceilf((float)(h + 1) / ss), wwend = (int)ceilf((float)(w + 1) / ss);
          hhstart = imin(imax(hhstart, 0), height_sp); hhend = imin(imax(hhend, 0), height_sp);
          wwstart = imin(imax(wwstart, 0), width_sp);  wwend = imin(imax(wwend, 0), width_sp);
          int prev = -1;
          for (int hh = hhstart; hh < hhend && !v; ++hh)
            for (int ww = wwstart; ww < wwend; ++ww) {
              int id = spp[hh * width_sp + ww];
              if (id == prev) continue;
              prev = id;
              if (ohl[id] == 1) { v = 1; break; }
            }
        }
        mois[h * width + w] = v;
      }
    int roi_width = imax(roi_end_w - roi_start_w + 1, 1);
    int roi_height = imax(roi_end_h - roi_start_h + 1, 1);
    float bin_h = (float)roi_height / (float)pooled_h;
    float bin_w = (float)roi_width / (float)pooled_w;
    for (int ph = 0; ph < pooled_h; ++ph)
      for (int pw = 0; pw < pooled_w; ++pw) {
        int hstart = (int)floorf((float)ph * bin_h), wstart = (int)floorf((float)pw * bin_w);
        int hend = (int)ceilf((float)(ph + 1) * bin_h), wend = (int)ceilf((float)(pw + 1) * bin_w);
        hstart = imin(imax(hstart + roi_start_h, 0), height);
        hend = imin(imax(hend + roi_start_h, 0), height);
        wstart = imin(imax(wstart + roi_start_w, 0), width);
        wend = imin(imax(wend + roi_start_w, 0), width);
        int is_empty = (hend <= hstart) || (wend <= wstart);
        for (int c = 0; c < channels; ++c) {
          const float* in = input + ((size_t)b * channels + c) * height * width;
          float maxval = is_empty ? 0.f : -FLT_MAX;
          int maxidx = -1;
          for (int h = hstart; h < hend; ++h)
            for (int w = wstart; w < wend; ++w) {
              int idx = h * width + w;
              if (mois[idx] == 0) continue;
              if (in[idx] > maxval) { maxval = in[idx]; maxidx = idx; }
            }
          size_t o = (((size_t)n * channels + c) * pooled_h + ph) * pooled_w + pw;
          output[o] = maxidx == -1 ? 0.f : maxval;
          argmax[o] = maxidx;
        }
      }
  }
  free(mois);
}

/* ------------------------------------------------------------------------
 * pairwise IoU + Matcher (OICR label assignment core).
 * detectron2/structures/boxes.py:345-391 (pairwise_intersection/pairwise_iou),
 * detectron2/modeling/matcher.py:61-103 with thresholds=[thr], labels=[0,1],
 * allow_low_quality_matches=False (roi_heads.py proposal_matcher).
 * gt [G,4], props [R,4] -> matched_idx[R] (first max over G), matched_val[R],
 * label[R] (1 iff val >= thr).  G == 0: idx 0, label 0 (matcher.py:79-91).
 * ------------------------------------------------------------------------ */
static float iou_pair(const float* a, const float* b) {
  /* boxes.py:360-366: wh = min(a[2:], b[2:]) - max(a[:2], b[:2]); clamp(min=0); prod */
  float w = fminf(a[2], b[2]) - fmaxf(a[0], b[0]);
  float h = fminf(a[3], b[3]) - fmaxf(a[1], b[1]);
  if (w < 0.f) w = 0.f;
  if (h < 0.f) h = 0.f;
  float inter = w * h;
  float area_a = (a[2] - a[0]) * (a[3] - a[1]); /* Boxes.area(), boxes.py:184-193 */
  float area_b = (b[2] - b[0]) * (b[3] - b[1]);
  /* boxes.py:385-389 */
  return inter > 0.f ? inter / (area_a + area_b - inter) : 0.f;
}

ORACLE_API void oracle_pairwise_iou(const float* boxes1, int n1, const float* boxes2, int n2,
                                    float* iou /* [n1,n2] */) {
  for (int i = 0; i < n1; ++i)
    for (int j = 0; j < n2; ++j) iou[(size_t)i * n2 + j] = iou_pair(boxes1 + 4 * i, boxes2 + 4 * j);
}

ORACLE_API void oracle_match(const float* gt, int num_gt, const float* props, int num_props,
                             float thr, int64_t* matched_idx, float* matched_val, int8_t* label) {
  for (int r = 0; r < num_props; ++r) {
    if (num_gt == 0) {
      matched_idx[r] = 0; matched_val[r] = 0.f; label[r] = 0;
      continue;
    }
    float best = -1.f;
    int bi = 0;
    for (int g = 0; g < num_gt; ++g) {
      float v = iou_pair(gt + 4 * g, props + 4 * r);
      if (v > best) { best = v; bi = g; } /* max(dim=0): first maximal index */
    }
    matched_idx[r] = bi;
    matched_val[r] = best;
    label[r] = best >= thr ? 1 : 0;
  }
}

/* ------------------------------------------------------------------------
 * NMS.  torchvision.ops.nms CPU kernel (third-party dependency of
 * detectron2/layers/nms.py:6-7,21,26; version unpinned by the reference,
 * torchvision 0.26 here): boxes visited in stable descending-score order; box j
 * is suppressed by a kept box i iff inter/(area_i + area_j - inter) > thr.
 * Returns the number kept; keep[] holds original indices in visiting order.
 * ------------------------------------------------------------------------ */
typedef struct { float s; int i; } ScoreIdx;
static int cmp_desc_stable(const void* a, const void* b) {
  const ScoreIdx* x = (const ScoreIdx*)a; const ScoreIdx* y = (const ScoreIdx*)b;
  if (x->s > y->s) return -1;
  if (x->s < y->s) return 1;
  return x->i - y->i;
}

ORACLE_API int oracle_nms(const float* boxes, const float* scores, int n, float thr, int64_t* keep) {
  ScoreIdx* order = (ScoreIdx*)malloc((size_t)(n ? n : 1) * sizeof(ScoreIdx));
  unsigned char* sup = (unsigned char*)calloc((size_t)(n ? n : 1), 1);
  for (int i = 0; i < n; ++i) { order[i].s = scores[i]; order[i].i = i; }
  qsort(order, (size_t)n, sizeof(ScoreIdx), cmp_desc_stable);
  int nk = 0;
  for (int a = 0; a < n; ++a) {
    int i = order[a].i;
    if (sup[i]) continue;
    keep[nk++] = i;
    const float* bi = boxes + 4 * (size_t)i;
    float iarea = (bi[2] - bi[0]) * (bi[3] - bi[1]);
    for (int c = a + 1; c < n; ++c) {
      int j = order[c].i;
      if (sup[j]) continue;
      const float* bj = boxes + 4 * (size_t)j;
      float xx1 = fmaxf(bi[0], bj[0]), yy1 = fmaxf(bi[1], bj[1]);
      float xx2 = fminf(bi[2], bj[2]), yy2 = fminf(bi[3], bj[3]);
      float w = fmaxf(0.f, xx2 - xx1), h = fmaxf(0.f, yy2 - yy1);
      float inter = w * h;
      float jarea = (bj[2] - bj[0]) * (bj[3] - bj[1]);
      float ovr = inter / (iarea + jarea - inter);
      if (ovr > thr) sup[j] = 1;
    }
  }
  free(order);
  free(sup);
  return nk;
}

/* ---------------------------------------------------------------------------------------------
 * PCL loss -- restatement of projects/WSL/wsl/layers/csrc/pcl_loss/pcl_loss_cpu.cpp
 *   forward  :8-57   (eps 1e-6, float accumulation in index order, logf of the clamped probability)
 *   backward :59-115 (eps 1e-5; top_grad unused)
 * Pinned bit for bit to the reference's own file built in place (oracle/_ref/libref_pclloss.so,
 * tests/test_ref_native.py). */
void oracle_pcl_loss_forward(const float* prob, const float* labels, const float* clw, const float* pc_labels,
                             const float* pc_probs, const float* iclw, const float* im_labels, int batch,
                             int channels, int num_positive, float* out) {
  const float eps = 1e-6f;
  for (int c = 0; c < channels; c++) {
    out[c] = 0;
    if (im_labels[c] != 0) {
      if (c == 0) {
        for (int i = 0; i < batch; i++)
          if (labels[i] == 0) out[c] -= clw[i] * logf(fmaxf(prob[i * channels + c], eps));
      } else {
        for (int i = 0; i < num_positive; i++)
          if (pc_labels[i] == c) out[c] -= iclw[i] * logf(fmaxf(pc_probs[i], eps));
      }
    }
  }
}

void oracle_pcl_loss_backward(const float* prob, const float* labels, const float* clw, const float* gt_assignment,
                              const float* pc_labels, const float* pc_probs, const float* pc_count,
                              const float* iclw, const float* im_labels, int batch, int channels, float* grad) {
  const float eps = 1e-5f;
  (void)pc_labels;
  for (int i = 0; i < batch; i++) {
    for (int c = 0; c < channels; c++) {
      grad[i * channels + c] = 0;
      if (im_labels[c] != 0) {
        if (c == 0) {
          if (labels[i] == 0) grad[i * channels + c] = -clw[i] / fmaxf(prob[i * channels + c], eps);
        } else if (labels[i] == c) {
          int p = (int)gt_assignment[i];
          grad[i * channels + c] = -iclw[p] / fmaxf(pc_count[p] * pc_probs[p], eps);
        }
      }
    }
  }
}
