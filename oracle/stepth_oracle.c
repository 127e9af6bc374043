/*
 * stepth_oracle.c -- CPU restatement of nikilark/stepth's hot path (see stepth_oracle.h).
 *
 * TEST INFRASTRUCTURE ONLY -- PARITY UNPINNED (no reference tests / golden vectors exist; the
 * reference cannot be compiled in this environment).  Every function cites the reference lines it
 * follows; loops keep the reference's order, including its duplicate probes, so that the timed CPU
 * baseline does the work the reference does.
 *
 * build: make -C oracle   (gcc -O2 -ffp-contract=off: Rust never contracts a*b+c into an FMA)
 */
#include "stepth_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * helpers.rs:3-7   distance_dot_dot: (((x1-x2)^2 + (y1-y2)^2) as f64).sqrt() as u32
 * ---------------------------------------------------------------------------------------------- */
uint32_t orc_distance_dot_dot(uint32_t x1, uint32_t y1, uint32_t x2, uint32_t y2) {
    int64_t dx = (int64_t)x1 - (int64_t)x2;
    int64_t dy = (int64_t)y1 - (int64_t)y2;
    double r = sqrt((double)(dx * dx + dy * dy));
    return (uint32_t)r; /* f64 -> u32 `as` cast: truncation toward zero; magnitudes here are tiny */
}

/* disage::pixels::PixelOpps<[u8;3]>: v.substract(what).lt(precision)   helpers.rs:35 (source absent) */
static inline int pixel_matches(const uint8_t *v, const uint8_t *what, const uint8_t *prec, int mode) {
    switch (mode) {
    case ORC_PIXELOPPS_ABS_ALL:
        for (int c = 0; c < 3; ++c) {
            int d = (int)v[c] - (int)what[c];
            if (d < 0) d = -d;
            if (!(d < (int)prec[c])) return 0;
        }
        return 1;
    case ORC_PIXELOPPS_SAT_ALL:
        for (int c = 0; c < 3; ++c) {
            int d = (int)v[c] - (int)what[c];
            if (d < 0) d = 0;
            if (!(d < (int)prec[c])) return 0;
        }
        return 1;
    default: {
        int s = 0, p = 0;
        for (int c = 0; c < 3; ++c) {
            int d = (int)v[c] - (int)what[c];
            s += d < 0 ? -d : d;
            p += prec[c];
        }
        return s < p;
    }
    }
}

/* ------------------------------------------------------------------------------------------------
 * helpers.rs:9-54   distance_dot_array
 *   for current_step in 0..max:                                   :26
 *     for (main, sub, order) in [(y, x, true), (x, y, false)]:    :28   rows pass, then columns pass
 *       for i in [main + step, main - step]:                      :29
 *         for j in sub - step ..= sub + step:                     :30
 *           (point_y, point_x) = order ? (i, j) : (j, i)          :31
 *           get2d -> None (out of bounds, negative wraps) => continue   :19-25,44
 *           Some(v): still_in_bounds = true; match => return (distance, pos)   :33-41
 *     if !still_in_bounds: break                                  :49-51
 * ---------------------------------------------------------------------------------------------- */
static int dda_impl(const uint8_t what[3], const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                    uint32_t from_x, uint32_t from_y, uint32_t max, const uint8_t precision[3],
                    int pixelopps, uint32_t *dist, uint32_t *pos_x, uint32_t *pos_y,
                    uint64_t *probes, uint64_t *unique) {
    const int64_t x = (int64_t)from_x, y = (int64_t)from_y;
    uint64_t n_probes = 0, n_unique = 0;
    for (int64_t step = 0; step < (int64_t)max; ++step) {
        int still_in_bounds = 0;
        for (int pass = 0; pass < 2; ++pass) {
            const int order = (pass == 0);
            const int64_t main_c = order ? y : x;
            const int64_t sub_c = order ? x : y;
            const int64_t is[2] = {main_c + step, main_c - step};
            for (int k = 0; k < 2; ++k) {
                const int64_t i = is[k];
                for (int64_t j = sub_c - step; j < sub_c + step + 1; ++j) {
                    const int64_t point_y = order ? i : j;
                    const int64_t point_x = order ? j : i;
                    /* array.get(i as usize): negative i64 wraps to a huge usize -> None */
                    if (point_y < 0 || point_y >= (int64_t)ha) continue;
                    if (point_x < 0 || point_x >= (int64_t)wa) continue;
                    const uint8_t *v = add_rgb + ((size_t)point_y * wa + (size_t)point_x) * 3;
                    still_in_bounds = 1;
                    ++n_probes;
                    /* a probe repeats an earlier one of the same ring when it is a corner of the column pass, or any
                     * probe of ring 0 after the first (bench.py's algorithmic-work figure counts each pixel once) */
                    if (step == 0 ? (pass == 0 && k == 0) : (order || (j != sub_c - step && j != sub_c + step)))
                        ++n_unique;
                    if (pixel_matches(v, what, precision, pixelopps)) {
                        if (dist) *dist = orc_distance_dot_dot(from_x, from_y, (uint32_t)point_x, (uint32_t)point_y);
                        if (pos_x) *pos_x = (uint32_t)point_x;
                        if (pos_y) *pos_y = (uint32_t)point_y;
                        if (probes) *probes = n_probes;
                        if (unique) *unique = n_unique;
                        return 1;
                    }
                }
            }
        }
        if (!still_in_bounds) break;
    }
    if (probes) *probes = n_probes;
    if (unique) *unique = n_unique;
    return 0;
}

int orc_distance_dot_array(const uint8_t what[3], const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                           uint32_t from_x, uint32_t from_y, uint32_t max, const uint8_t precision[3],
                           int pixelopps, uint32_t *dist, uint32_t *pos_x, uint32_t *pos_y,
                           uint64_t *probes) {
    return dda_impl(what, add_rgb, wa, ha, from_x, from_y, max, precision, pixelopps, dist, pos_x, pos_y, probes, NULL);
}

/* ------------------------------------------------------------------------------------------------
 * depth_image.rs:110-122   the per-block loop, with rayon's par_chunks_mut(len / 8) restated as
 * one pthread per chunk (8 or 9 chunks) when threads == 0.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    /* block mode */
    const orc_block *blocks;
    /* dense mode */
    const uint8_t *main_px;
    uint32_t main_stride, w, h;
    const uint32_t *indices; /* optional subset (dense) */
    /* common */
    const uint8_t *add_rgb;
    uint32_t wa, ha, max_rings;
    uint8_t precision[3];
    int pixelopps;
    size_t begin, end;
    uint8_t *d8_out;
    uint64_t probes, unique;
} search_job;

static void *search_worker(void *arg) {
    search_job *jb = (search_job *)arg;
    uint64_t total = 0, total_u = 0;
    for (size_t i = jb->begin; i < jb->end; ++i) {
        uint8_t what[3];
        uint32_t mx, my;
        if (jb->blocks) {
            const orc_block *p = &jb->blocks[i];
            /* depth_image.rs:114-117  middle = ((pos.x + size.width)/2, (pos.y + size.height)/2)  (literal) */
            mx = (p->x + p->w) / 2;
            my = (p->y + p->h) / 2;
            what[0] = p->value[0]; what[1] = p->value[1]; what[2] = p->value[2];
        } else {
            size_t q = jb->indices ? jb->indices[i] : i;
            uint32_t px = (uint32_t)(q % jb->w), py = (uint32_t)(q / jb->w);
            mx = (px + 1) / 2; /* 1x1 block: (x + 1)/2 */
            my = (py + 1) / 2;
            const uint8_t *m = jb->main_px + q * jb->main_stride;
            what[0] = m[0]; what[1] = m[1]; what[2] = m[2];
        }
        uint32_t d = 0; /* .unwrap_or((u32::MIN, Position::new(0,0)))   depth_image.rs:120 */
        uint64_t pr = 0, un = 0;
        dda_impl(what, jb->add_rgb, jb->wa, jb->ha, mx, my, jb->max_rings, jb->precision,
                 jb->pixelopps, &d, NULL, NULL, &pr, &un);
        total += pr;
        total_u += un;
        jb->d8_out[i] = (uint8_t)d; /* `d as u8`: wraps mod 256   depth_image.rs:121 */
    }
    jb->probes = total;
    jb->unique = total_u;
    return NULL;
}

static uint64_t g_last_unique; /* unique probes of the last run_search (read by orc_last_unique_probes; bench bookkeeping) */

uint64_t orc_last_unique_probes(void) { return g_last_unique; }

static int32_t run_search(search_job proto, size_t n, int threads, uint64_t *probes_total) {
    if (n == 0) return ORC_TOO_FEW_BLOCKS;
    size_t chunk;
    if (threads <= 0) {
        chunk = n / 8; /* depth_image.rs:111 */
        if (chunk == 0) return ORC_TOO_FEW_BLOCKS; /* par_chunks_mut(0) panics   :112 */
    } else {
        chunk = (n + (size_t)threads - 1) / (size_t)threads;
    }
    size_t n_chunks = (n + chunk - 1) / chunk;
    search_job *jobs = (search_job *)malloc(n_chunks * sizeof(search_job));
    pthread_t *tids = (pthread_t *)malloc(n_chunks * sizeof(pthread_t));
    if (!jobs || !tids) { free(jobs); free(tids); return ORC_INVALID_ARG; }
    for (size_t c = 0; c < n_chunks; ++c) {
        jobs[c] = proto;
        jobs[c].begin = c * chunk;
        jobs[c].end = (c + 1) * chunk < n ? (c + 1) * chunk : n;
        jobs[c].probes = 0;
    }
    if (n_chunks == 1) {
        search_worker(&jobs[0]);
    } else {
        for (size_t c = 0; c < n_chunks; ++c) pthread_create(&tids[c], NULL, search_worker, &jobs[c]);
        for (size_t c = 0; c < n_chunks; ++c) pthread_join(tids[c], NULL);
    }
    if (probes_total) {
        uint64_t t = 0;
        for (size_t c = 0; c < n_chunks; ++c) t += jobs[c].probes;
        *probes_total = t;
    }
    {
        uint64_t t = 0;
        for (size_t c = 0; c < n_chunks; ++c) t += jobs[c].unique;
        g_last_unique = t;
    }
    free(jobs);
    free(tids);
    return ORC_OK;
}

int32_t orc_search_blocks(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const orc_block *blocks,
                          size_t n_blocks, const uint8_t precision[3], uint32_t max_rings, int pixelopps,
                          int threads, uint8_t *d8_out, uint64_t *probes_total) {
    if (!add_rgb || !blocks || !d8_out || !precision) return ORC_INVALID_ARG;
    search_job jb;
    memset(&jb, 0, sizeof jb);
    jb.blocks = blocks;
    jb.add_rgb = add_rgb; jb.wa = wa; jb.ha = ha; jb.max_rings = max_rings;
    memcpy(jb.precision, precision, 3);
    jb.pixelopps = pixelopps;
    jb.d8_out = d8_out;
    return run_search(jb, n_blocks, threads, probes_total);
}

int32_t orc_search_dense(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                         const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                         uint32_t max_rings, int pixelopps, int threads, uint8_t *d8_out,
                         uint64_t *probes_total) {
    if (!main_px || !add_rgb || !d8_out || !precision || (main_stride != 3 && main_stride != 4))
        return ORC_INVALID_ARG;
    search_job jb;
    memset(&jb, 0, sizeof jb);
    jb.main_px = main_px; jb.main_stride = main_stride; jb.w = w; jb.h = h;
    jb.add_rgb = add_rgb; jb.wa = wa; jb.ha = ha; jb.max_rings = max_rings;
    memcpy(jb.precision, precision, 3);
    jb.pixelopps = pixelopps;
    jb.d8_out = d8_out;
    return run_search(jb, (size_t)w * h, threads, probes_total);
}

int32_t orc_search_dense_subset(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                                const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                                uint32_t max_rings, int pixelopps, int threads, const uint32_t *indices,
                                size_t n_idx, uint8_t *d8_out, uint64_t *probes_total) {
    if (!main_px || !add_rgb || !d8_out || !precision || !indices || (main_stride != 3 && main_stride != 4))
        return ORC_INVALID_ARG;
    search_job jb;
    memset(&jb, 0, sizeof jb);
    jb.main_px = main_px; jb.main_stride = main_stride; jb.w = w; jb.h = h;
    jb.indices = indices;
    jb.add_rgb = add_rgb; jb.wa = wa; jb.ha = ha; jb.max_rings = max_rings;
    memcpy(jb.precision, precision, 3);
    jb.pixelopps = pixelopps;
    jb.d8_out = d8_out;
    return run_search(jb, n_idx, threads, probes_total);
}

/* ------------------------------------------------------------------------------------------------
 * depth_image.rs:124-129
 *   max = pixels.iter().max_by_key(|p| p.value[0]).value[0]
 *   value = (value as u64 * u8::MAX as u64 / max as u64) as u8        (max == 0 -> panic)
 * ---------------------------------------------------------------------------------------------- */
int32_t orc_normalise(uint8_t *d8, size_t n, uint8_t *max_out) {
    uint8_t mx = 0;
    for (size_t i = 0; i < n; ++i)
        if (d8[i] > mx) mx = d8[i];
    if (max_out) *max_out = mx;
    if (mx == 0) return ORC_DIV_BY_ZERO_MAX;
    for (size_t i = 0; i < n; ++i) d8[i] = (uint8_t)((uint64_t)d8[i] * 255u / (uint64_t)mx);
    return ORC_OK;
}

/* depth_image.rs:130-132: DiscreteImage::collect() fills each block's rectangle with its value
 * (blocks are assumed to tile the image: SURVEY Appendix B); luma of (v,v,v) is v. */
int32_t orc_paint_blocks(const orc_block *blocks, const uint8_t *values, size_t n_blocks, uint32_t w,
                         uint32_t h, uint8_t *luma_out) {
    for (size_t b = 0; b < n_blocks; ++b) {
        const orc_block *p = &blocks[b];
        if ((uint64_t)p->x + p->w > w || (uint64_t)p->y + p->h > h) return ORC_INVALID_ARG;
        for (uint32_t yy = p->y; yy < p->y + p->h; ++yy)
            memset(luma_out + (size_t)yy * w + p->x, values[b], p->w);
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------------
 * depth_image.rs:133-134   DynamicImage::resize(w, h, Gaussian) at unit scale.
 * image 0.23.14 (Cargo.lock:171-174; source not under /root/reference) imageops/sample.rs:
 *   resize() = horizontal_sample(&vertical_sample(image, nheight, filter), nwidth, filter)
 *   gaussian(x, r) = ((2*PI).sqrt() * r).recip() * (-x.powi(2) / (2.0 * r.powi(2))).exp(),  r = 0.5, support 3.0
 *   per output o:  in = (o + 0.5) * ratio;  left = clamp(floor(in - support), 0, n-1);
 *                  right = clamp(ceil(in + support), left+1, n);  in -= 0.5;
 *                  w_i = kernel((i - in) / sratio), sum += w_i;  w_i /= sum;
 *                  t += px_i as f32 * w_i  (separate mul and add);  out = round(clamp(t, 0, 255))
 * The generic sampler of 0.23.x returns ImageBuffer<P, Vec<S>> from vertical_sample, i.e. a *rounded u8*
 * intermediate (ORC_RESAMPLE_U8, default); the f32 intermediate arrived with Rgba32FImage in 0.24
 * (ORC_RESAMPLE_F32, kept as a switch because the crate source is not available to confirm).
 * ---------------------------------------------------------------------------------------------- */
static float gaussian_kernel(float x) {
    const float r = 0.5f;
    float norm = 1.0f / (sqrtf(2.0f * 3.14159265358979323846f) * r);
    return norm * expf(-(x * x) / (2.0f * (r * r)));
}

/* taps for output index o along an axis of n samples at ratio 1 */
static int gaussian_taps(uint32_t o, uint32_t n, uint32_t *left_out, float ws[8]) {
    const float ratio = 1.0f, sratio = 1.0f, support = 3.0f * sratio;
    float inputx = ((float)o + 0.5f) * ratio;
    int64_t left = (int64_t)floorf(inputx - support);
    if (left < 0) left = 0;
    if (left > (int64_t)n - 1) left = (int64_t)n - 1;
    int64_t right = (int64_t)ceilf(inputx + support);
    if (right < left + 1) right = left + 1;
    if (right > (int64_t)n) right = (int64_t)n;
    inputx = inputx - 0.5f;
    int cnt = 0;
    float sum = 0.0f;
    for (int64_t i = left; i < right; ++i) {
        float wgt = gaussian_kernel(((float)i - inputx) / sratio);
        ws[cnt++] = wgt;
        sum += wgt;
    }
    for (int k = 0; k < cnt; ++k) ws[k] /= sum;
    *left_out = (uint32_t)left;
    return cnt;
}

static inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

int32_t orc_gaussian_resample_unit(const uint8_t *src, uint32_t w, uint32_t h, int mode, uint8_t *dst) {
    if (!src || !dst || w == 0 || h == 0) return ORC_INVALID_ARG;
    size_t n = (size_t)w * h;
    if (mode == ORC_RESAMPLE_OFF) { memmove(dst, src, n); return ORC_OK; }
    float *tmp = (float *)malloc(n * sizeof(float));
    if (!tmp) return ORC_INVALID_ARG;
    float ws[8];
    uint32_t left;
    /* vertical_sample: for outy { weights; for x { t = sum_i px(x, left+i) * w_i } } */
    for (uint32_t oy = 0; oy < h; ++oy) {
        int cnt = gaussian_taps(oy, h, &left, ws);
        for (uint32_t x = 0; x < w; ++x) {
            float t = 0.0f;
            for (int k = 0; k < cnt; ++k) {
                float v = (float)src[(size_t)(left + (uint32_t)k) * w + x];
                t += v * ws[k];
            }
            if (mode == ORC_RESAMPLE_U8) t = roundf(clampf(t, 0.0f, 255.0f)); /* FloatNearest -> u8 */
            tmp[(size_t)oy * w + x] = t;
        }
    }
    /* horizontal_sample: for outx { weights; for y { ... } } */
    for (uint32_t ox = 0; ox < w; ++ox) {
        int cnt = gaussian_taps(ox, w, &left, ws);
        for (uint32_t y = 0; y < h; ++y) {
            float t = 0.0f;
            for (int k = 0; k < cnt; ++k) t += tmp[(size_t)y * w + left + (uint32_t)k] * ws[k];
            dst[(size_t)y * w + ox] = (uint8_t)roundf(clampf(t, 0.0f, 255.0f));
        }
    }
    free(tmp);
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------------
 * depth_image.rs:91-136   load_depth_from_additional, steps 5-11 (blocks are the input: the disage
 * discretiser that produces them is host-side and its source is absent).
 * ---------------------------------------------------------------------------------------------- */
int32_t orc_depth_from_blocks(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const orc_block *blocks,
                              size_t n_blocks, const uint8_t precision[3], uint32_t max_rings, uint32_t w,
                              uint32_t h, int pixelopps, int resample, int threads, uint8_t *depth_out,
                              uint8_t *raw_d8_out) {
    if (!depth_out) return ORC_INVALID_ARG;
    uint8_t *d8 = (uint8_t *)malloc(n_blocks ? n_blocks : 1);
    size_t n_px = (size_t)w * h;
    uint8_t *painted = (uint8_t *)calloc(n_px > 0 ? n_px : 1, 1);
    int32_t st = orc_search_blocks(add_rgb, wa, ha, blocks, n_blocks, precision, max_rings, pixelopps, threads, d8, NULL);
    if (st == ORC_OK && raw_d8_out) memcpy(raw_d8_out, d8, n_blocks);
    if (st == ORC_OK) st = orc_normalise(d8, n_blocks, NULL);
    if (st == ORC_OK) st = orc_paint_blocks(blocks, d8, n_blocks, w, h, painted);
    if (st == ORC_OK) st = orc_gaussian_resample_unit(painted, w, h, resample, depth_out);
    free(d8);
    free(painted);
    return st;
}

int32_t orc_depth_dense(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                        const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                        uint32_t max_rings, int pixelopps, int resample, int threads, uint8_t *depth_out,
                        uint8_t *raw_d8_out) {
    if (!depth_out) return ORC_INVALID_ARG;
    size_t n = (size_t)w * h;
    uint8_t *d8 = (uint8_t *)malloc(n ? n : 1);
    int32_t st = orc_search_dense(main_px, main_stride, w, h, add_rgb, wa, ha, precision, max_rings, pixelopps, threads, d8, NULL);
    if (st == ORC_OK && raw_d8_out) memcpy(raw_d8_out, d8, n);
    if (st == ORC_OK) st = orc_normalise(d8, n, NULL);
    if (st == ORC_OK) st = orc_gaussian_resample_unit(d8, w, h, resample, depth_out);
    free(d8);
    return st;
}

/* depth_image.rs:225-227 */
void orc_invert_depth(uint8_t *depth, size_t n) {
    for (size_t i = 0; i < n; ++i) depth[i] = (uint8_t)(255 - depth[i]);
}

/* ------------------------------------------------------------------------------------------------
 * depth_image.rs:162-218   depth_split: 1-D Lloyd iteration over the depth bytes, restated literally
 * (one pass over the whole array per iteration, like the reference's `inner`).
 *   clusters: HashMap<u8, Vec<u8>> keyed by centre value (duplicate centres collapse)      :166-170
 *   assignment: prev_centers.iter().min_by(|x,y| |x-item| cmp |y-item|) -> first minimum   :171-184
 *   new centroid = sum / len.max(1) as u8, sorted                                          :185-189
 *   converged iff zip(prev, new) all equal (zip truncates)                                 :190-196
 *   result per new centroid c: clusters.get(c).unwrap() -> (min, max); unwrap panics       :197-207
 * ---------------------------------------------------------------------------------------------- */
static int cmp_u8(const void *a, const void *b) { return (int)*(const uint8_t *)a - (int)*(const uint8_t *)b; }

int32_t orc_depth_split(const uint8_t *depth, size_t n, uint8_t zones, uint8_t *ranges_out, uint32_t *n_out,
                        int *has_values_out, uint8_t *centroids_out, uint32_t *iters_out) {
    if (!ranges_out || !n_out || !has_values_out) return ORC_INVALID_ARG;
    if (iters_out) *iters_out = 0;
    if (zones < 2) { /* :163-165  vec![(None, None)] */
        *n_out = 1;
        *has_values_out = 0;
        return ORC_OK;
    }
    *has_values_out = 1;
    if (n == 0) return ORC_DEGENERATE_DEPTH; /* .iter().min().unwrap() on empty   :212 */
    uint8_t img_min = 255, img_max = 0;
    for (size_t i = 0; i < n; ++i) {
        if (depth[i] < img_min) img_min = depth[i];
        if (depth[i] > img_max) img_max = depth[i];
    }
    /* :214-216  (img_min..img_max).step_by(((img_max - img_min) / (zones - 1)) as usize - 1) */
    uint32_t q = (uint32_t)((uint8_t)(img_max - img_min) / (uint8_t)(zones - 1));
    if (q == 0) return ORC_DEGENERATE_DEPTH; /* usize underflow (debug) / empty centres then unwrap(None) */
    uint32_t step = q - 1;
    if (step == 0) return ORC_DEGENERATE_DEPTH; /* step_by(0): "assertion failed: step != 0" */
    uint8_t prev[256];
    uint32_t n_prev = 0;
    for (uint32_t c = img_min; c < img_max; c += step) prev[n_prev++] = (uint8_t)c;

    for (uint32_t iter = 1; iter <= 100000; ++iter) {
        /* clusters keyed by centre value */
        uint64_t sum[256];
        uint64_t cnt[256];
        uint8_t cmin[256], cmax[256];
        uint8_t is_key[256];
        memset(sum, 0, sizeof sum);
        memset(cnt, 0, sizeof cnt);
        memset(cmin, 255, sizeof cmin);
        memset(cmax, 0, sizeof cmax);
        memset(is_key, 0, sizeof is_key);
        for (uint32_t k = 0; k < n_prev; ++k) is_key[prev[k]] = 1;
        if (n_prev == 0) return ORC_DEGENERATE_DEPTH; /* min_by on empty -> unwrap(None) */
        for (size_t i = 0; i < n; ++i) {
            int item = depth[i];
            int best = 0, best_d = abs((int)prev[0] - item);
            for (uint32_t k = 1; k < n_prev; ++k) {
                int d = abs((int)prev[k] - item);
                if (d < best_d) { best_d = d; best = (int)k; } /* strict: first minimum wins */
            }
            uint8_t key = prev[best];
            sum[key] += (uint64_t)item;
            cnt[key] += 1;
            if (depth[i] < cmin[key]) cmin[key] = depth[i];
            if (depth[i] > cmax[key]) cmax[key] = depth[i];
        }
        /* new centroids (one per distinct key), sorted; remember which cluster each came from */
        uint8_t newc[256];
        uint32_t n_new = 0;
        for (int key = 0; key < 256; ++key)
            if (is_key[key]) newc[n_new++] = (uint8_t)(sum[key] / (cnt[key] ? cnt[key] : 1));
        qsort(newc, n_new, 1, cmp_u8);
        int same = 1;
        uint32_t m = n_prev < n_new ? n_prev : n_new;
        for (uint32_t k = 0; k < m; ++k)
            if (prev[k] != newc[k]) { same = 0; break; }
        if (same) {
            /* :197-207  clusters.get(c).unwrap(): c must be a key, and its Vec must be non-empty */
            for (uint32_t k = 0; k < n_new; ++k) {
                uint8_t c = newc[k];
                if (!is_key[c] || cnt[c] == 0) return ORC_DEGENERATE_DEPTH;
                ranges_out[2 * k] = cmin[c];
                ranges_out[2 * k + 1] = cmax[c];
                if (centroids_out) centroids_out[k] = c;
            }
            *n_out = n_new;
            if (iters_out) *iters_out = iter;
            return ORC_OK;
        }
        memcpy(prev, newc, n_new);
        n_prev = n_new;
    }
    return ORC_NO_CONVERGENCE;
}

/* depth_image.rs:229-245   mask = MASK_TRUE; MASK_FALSE where d < from || d > to */
void orc_slice(const uint8_t *depth, size_t n, int from, int to, uint8_t *mask_out) {
    uint8_t f = from < 0 ? 0 : (uint8_t)from;   /* from.unwrap_or(0) */
    uint8_t t = to < 0 ? 255 : (uint8_t)to;     /* to.unwrap_or(u8::MAX) */
    for (size_t i = 0; i < n; ++i) mask_out[i] = (depth[i] < f || depth[i] > t) ? 0 : 255;
}

/* depth_image.rs:220-223   let (from, to) = self.depth_split(2)[0]; self.slice(from, to) */
int32_t orc_select_foreground(const uint8_t *depth, size_t n, uint8_t *mask_out, uint8_t from_to_out[2]) {
    uint8_t ranges[512];
    uint32_t n_out = 0;
    int has = 0;
    int32_t st = orc_depth_split(depth, n, 2, ranges, &n_out, &has, NULL, NULL);
    if (st != ORC_OK) return st;
    if (n_out == 0) return ORC_DEGENERATE_DEPTH; /* [0] on an empty Vec */
    if (from_to_out) { from_to_out[0] = ranges[0]; from_to_out[1] = ranges[1]; }
    orc_slice(depth, n, ranges[0], ranges[1], mask_out);
    return ORC_OK;
}

/* mask_image.rs:205-213   pixel = [0;4] where mask == MASK_FALSE (exactly 0) */
void orc_apply_mask(uint8_t *rgba, const uint8_t *mask, size_t n) {
    for (size_t i = 0; i < n; ++i)
        if (mask[i] == 0) memset(rgba + 4 * i, 0, 4);
}

/* Rust `f64 as u16`: saturating, NaN -> 0 */
static inline uint16_t f64_as_u16(double v) {
    if (!(v == v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 65535.0) return 65535;
    return (uint16_t)v;
}

/* operations.rs:3-36 */
void orc_normalize_brightness_luma16(const uint16_t *img1, size_t n1, const uint16_t *img2, size_t n2,
                                     double percent, uint16_t *out) {
    uint64_t fbr = 0, sbr = 0;
    for (size_t i = 0; i < n1; ++i) fbr += img1[i];
    for (size_t i = 0; i < n2; ++i) sbr += img2[i];
    fbr = fbr / (uint64_t)n1; /* :25  (n == 0 panics in the reference; callers must not pass it) */
    sbr = sbr / (uint64_t)n2;
    double diff = (double)sbr / (double)fbr;
    if (fabs(1.0 - diff) < percent) { memmove(out, img1, n1 * sizeof(uint16_t)); return; }
    for (size_t i = 0; i < n1; ++i) out[i] = f64_as_u16((double)img1[i] * diff);
}

/* operations.rs:38-80   note `img.len()` is the SUBPIXEL count (3 * pixels): the means are divided by it */
void orc_normalize_brightness_rgb16(const uint16_t *img1, size_t n1_px, const uint16_t *img2, size_t n2_px,
                                    double percent, uint16_t *out) {
    double rgb1[3] = {0, 0, 0}, rgb2[3] = {0, 0, 0};
    for (size_t i = 0; i < n1_px; ++i)
        for (int c = 0; c < 3; ++c) rgb1[c] += (double)img1[3 * i + c];
    for (size_t i = 0; i < n2_px; ++i)
        for (int c = 0; c < 3; ++c) rgb2[c] += (double)img2[3 * i + c];
    for (int c = 0; c < 3; ++c) {
        rgb1[c] /= (double)(3 * n1_px);
        rgb2[c] /= (double)(3 * n2_px);
    }
    double diff[3] = {rgb2[0] / rgb1[0], rgb2[1] / rgb1[1], rgb2[2] / rgb1[2]};
    if (fabs(1.0 - diff[0]) < percent && fabs(1.0 - diff[1]) < percent && fabs(1.0 - diff[2]) < percent) {
        memmove(out, img1, 3 * n1_px * sizeof(uint16_t));
        return;
    }
    for (size_t i = 0; i < n1_px; ++i)
        for (int c = 0; c < 3; ++c) out[3 * i + c] = f64_as_u16((double)img1[3 * i + c] * diff[c]);
}

/* ---- editor operations ------------------------------------------------------------------ */

/* `(v as f32 * m).max(0.0).min(255.0) as u8`: f32 product, clamp, truncation */
static uint8_t adjust_u8(uint8_t v, float m) {
    float t = (float)v * m;
    if (!(t > 0.0f)) t = 0.0f; /* f32::max(0.0) also maps NaN to 0.0 */
    if (t > 255.0f) t = 255.0f;
    return (uint8_t)t;
}

/* depth_image.rs:51-63: multiplier = d as f32 / 255.0 * 2.0, applied to r, g, b; alpha untouched */
void orc_highlight_depth(const uint8_t *rgba_in, const uint8_t *depth, size_t n, uint8_t *rgba_out) {
    for (size_t i = 0; i < n; ++i) {
        volatile float q = (float)depth[i] / 255.0f; /* volatile: keep the two roundings of the reference */
        const float m = q * 2.0f;
        rgba_out[4 * i + 0] = adjust_u8(rgba_in[4 * i + 0], m);
        rgba_out[4 * i + 1] = adjust_u8(rgba_in[4 * i + 1], m);
        rgba_out[4 * i + 2] = adjust_u8(rgba_in[4 * i + 2], m);
        rgba_out[4 * i + 3] = rgba_in[4 * i + 3];
    }
}

/* mask_image.rs:57-73: where mask == MASK_TRUE: r * 2.0, g * (1.0 / 2.0), b * (1.0 / 2.0) */
void orc_highlight_mask(const uint8_t *rgba_in, const uint8_t *mask, size_t n, uint8_t *rgba_out) {
    const float multiplier = 2.0f;
    for (size_t i = 0; i < n; ++i) {
        uint8_t r = rgba_in[4 * i], g = rgba_in[4 * i + 1], b = rgba_in[4 * i + 2];
        if (mask[i] == 255) {
            r = adjust_u8(r, multiplier);
            g = adjust_u8(g, 1.0f / multiplier);
            b = adjust_u8(b, 1.0f / multiplier);
        }
        rgba_out[4 * i] = r; rgba_out[4 * i + 1] = g; rgba_out[4 * i + 2] = b; rgba_out[4 * i + 3] = rgba_in[4 * i + 3];
    }
}

/* mask_image.rs:147-169 / 171-193, the equal-dimensions branch */
void orc_mask_and(uint8_t *mask, const uint8_t *other, size_t n) {
    for (size_t i = 0; i < n; ++i) mask[i] = (mask[i] == 255 && other[i] == 255) ? 255 : 0;
}
void orc_mask_or(uint8_t *mask, const uint8_t *other, size_t n) {
    for (size_t i = 0; i < n; ++i) mask[i] = (mask[i] == 255 || other[i] == 255) ? 255 : 0;
}
/* mask_image.rs:195-197 */
void orc_mask_not(uint8_t *mask, size_t n) {
    for (size_t i = 0; i < n; ++i) mask[i] = (uint8_t)(255 - mask[i]);
}

/* mask_image.rs:99-109: note `other.image.get_pixel(x, y)` uses the coordinates of SELF (literal) */
int32_t orc_image_replace(uint8_t *rgba, const uint8_t *mask, uint32_t w, uint32_t h, const uint8_t *other_rgba,
                          uint32_t ow, uint32_t oh, uint32_t start_x, uint32_t start_y) {
    uint64_t y1 = (uint64_t)start_y + oh, x1 = (uint64_t)start_x + ow;
    if (y1 > h) y1 = h;
    if (x1 > w) x1 = w;
    for (uint64_t y = start_y; y < y1; ++y)
        for (uint64_t x = start_x; x < x1; ++x) {
            if (mask[y * w + x] != 255) continue;
            if (x >= ow || y >= oh) return ORC_OUT_OF_BOUNDS; /* get_pixel panics */
            memcpy(rgba + (y * w + x) * 4, other_rgba + (y * ow + x) * 4, 4);
        }
    return ORC_OK;
}
