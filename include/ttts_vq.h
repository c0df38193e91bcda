/*
 * ttts_vq.h -- C ABI of the step directly behind the log-mel path: the DVAE codebook search, same
 * library (libttts_mel.so), same conventions as ttts_mel.h: plain pointers and sizes, caller-owned device buffers and
 * workspace, `stream` is a cudaStream_t passed as void*, 0 / negative TTTS_MEL_E* return codes, ttts_mel_last_error().
 *
 * Reference: `Quantize.forward` in eval mode (ttts/vqvae/xtts_dvae.py:72-127), as called by
 * `DiscreteVAE.get_codebook_indices` (xtts_dvae.py:327-334) from ttts/prepare/extract_vq.py:24-29:
 *     flatten = input.reshape(-1, dim)                                               (xtts_dvae.py:87)
 *     dist = flatten.pow(2).sum(1, keepdim=True) - 2 * flatten @ embed + embed.pow(2).sum(0, keepdim=True)   (:88)
 *     _, embed_ind = (-dist).max(1)                                                  (:89-90)   ttts_vq_nearest
 *     quantize = F.embedding(embed_ind, embed.transpose(0, 1))                       (:93, :129-130)   ttts_vq_embed_code
 *     diff = (quantize.detach() - input).pow(2).mean()                               (:119)
 *     quantize = input + (quantize - input).detach()                                 (:120)
 * The encoder in front of it (a convolutional model) and the EMA codebook update of training mode (:104-117) are not
 * part of this path.
 */
#ifndef TTTS_VQ_H_
#define TTTS_VQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ttts_vq_plan ttts_vq_plan;

/* The plan holds the codebook in the layouts the kernels read (tensor-core operand tiles, split hi / lo; row-major
 * [n_embed, dim] for the gather; |e_j|^2).  It replaces the `embed` buffer of the module (xtts_dvae.py:67-68).
 *   embed_host [dim, n_embed] float32, row-major: exactly the reference's buffer
 *   dim        1 .. 4096 (config.json: codebook_dim 512); n_embed 16 .. 2^24 (config.json: num_tokens 8192) */
int ttts_vq_plan_create(const float* embed_host, int32_t dim, int32_t n_embed, ttts_vq_plan** out_plan);
void ttts_vq_plan_destroy(ttts_vq_plan* plan);

/* Workspace for one ttts_vq_nearest call over n_rows vectors. */
size_t ttts_vq_workspace_bytes(const ttts_vq_plan* plan, int64_t n_rows);

/* embed_ind[i] = argmin_j |z_i - e_j|^2 (the lowest j among equal distances, as torch.max returns the first maximum),
 * evaluated as |e_j|^2 - 2 z_i . e_j with fp32-grade products (xtts_dvae.py:87-90).  Rows that contain a NaN get
 * index 0 (every distance is NaN; torch returns the first).
 *   z_dev           [n_rows, row_stride] float32, the first `dim` of every row are used
 *   embed_ind_dev   [n_rows] int64
 *   quantize_dev_or_null  [n_rows, dim] float32: z + (e[embed_ind] - z), the straight-through value of :120
 *   diff_dev_or_null      one float32: mean over n_rows * dim of (e[embed_ind] - z)^2   (:119)
 * Asynchronous on `stream`; no allocation, no synchronisation. */
int ttts_vq_nearest(const ttts_vq_plan* plan, const float* z_dev, int64_t row_stride, int64_t n_rows,
                    int64_t* embed_ind_dev, float* quantize_dev_or_null, float* diff_dev_or_null,
                    void* workspace_dev, size_t workspace_bytes, void* stream);

/* The same search for the encoder's own output layout [B, dim, T] (channel-major), i.e. `get_codebook_indices`
 * without the `permute((0, 2, 1))` + `reshape` copy of xtts_dvae.py:331,87: row n = b T + t.  Indices only.
 *   z_dev [B, dim, T] float32 with element (b, k, t) at b * batch_stride + k * chan_stride + t
 *   embed_ind_dev [B * T] int64 (= [B, T]) */
int ttts_vq_nearest_bct(const ttts_vq_plan* plan, const float* z_dev, int64_t batch_stride, int64_t chan_stride,
                        int64_t B, int64_t T, int64_t* embed_ind_dev,
                        void* workspace_dev, size_t workspace_bytes, void* stream);

/* `Quantize.embed_code` (xtts_dvae.py:129-130): out[i, :] = embed[:, embed_ind[i]].  Indices outside
 * [0, n_embed) raise the device status (checked by ttts_vq_status) and produce zeros. */
int ttts_vq_embed_code(const ttts_vq_plan* plan, const int64_t* embed_ind_dev, int64_t n_rows,
                       float* out_dev, int64_t out_row_stride, void* stream);

/* Synchronises `stream` and reports what the kernels of this plan flagged (pipeline time-out codes, bad indices);
 * 0 = clean.  Optional: for tests and for callers that want device-side errors surfaced as return codes. */
int ttts_vq_status(const ttts_vq_plan* plan, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* TTTS_VQ_H_ */
