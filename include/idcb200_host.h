/* idc_models_b200 — C ABI of the host-only library (idc_models_b200/_native/libidcb200_host.so, the sources under csrc/host).
 *
 * No CUDA dependency: loadable on a CPU-only box. Everything the reference gets from native code inside
 * TensorFlow / h5py / phe on the HOST side lives here: the tf.data runtime (shuffle buffer, batching,
 * prefetch), PNG decoding, the HDF5 subset Keras weight files need, the checkpoint container, the client scheduler of
 * the federated loops and the CPU restatement of the secure-aggregation arithmetic.
 *
 * Conventions: functions returning int return 0 on success and non-zero on failure; functions returning a handle
 * return NULL on failure; idcbh_last_error() then holds the reason (per calling thread). Handles are not thread
 * safe unless stated. All sizes are in elements unless the name says bytes.
 */
#ifndef IDCB200_HOST_H
#define IDCB200_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* idcbh_last_error(void);
int idcbh_version(void);

/* ---- input pipeline runtime: cache -> shuffle(buffer) -> batch -> prefetch -------------------------------------
 * (prepare_for_training, /root/reference/dist_model_tf_mobile.py:47-65). `images` [n][sample_floats] and `labels`
 * [n] stay owned by the caller. Worker threads assemble shuffled batches (tf.data's buffer-shuffle algorithm, a new
 * order every epoch) into the caller's ring of `n_slots` slots (pinned memory, so an asynchronous H2D copy can start
 * straight out of a slot). Batch k always lands in slot k % n_slots; a slot is refilled only after release(). */
void* idcbh_loader_create(const float* images, const float* labels, long long n, long long sample_floats, int batch,
                          int shuffle_buffer, int drop_remainder, int n_slots, float* const* slot_x,
                          float* const* slot_y, int n_threads, unsigned long long seed);
int idcbh_loader_start_epoch(void* h);                 /* no slot of the previous epoch may still be held */
int idcbh_loader_next(void* h, int* slot, int* count);  /* blocks; 1 = batch ready, 0 = end of epoch, < 0 = error */
int idcbh_loader_release(void* h, int slot);
long long idcbh_loader_num_batches(void* h);
void idcbh_loader_destroy(void* h);

/* ---- PNG stage: decode_png(channels=3) -> float32 in [0,1] -> bilinear resize (half-pixel centres) --------------
 * (/root/reference/dist_model_tf_mobile.py:37-45). Own inflate; every non-interlaced colour type of 8 bits or less.
 * Status / return: 0 decoded, 1 a valid PNG this decoder leaves to a general one (16-bit samples, Adam7), 2 corrupt
 * or unreadable. out is [H][W][3]; decode_files fills out[n][H][W][3] with `threads` workers and returns the number
 * of files whose status is not 0 (their rows of `out` are untouched). */
int idcbh_png_decode_f32(const uint8_t* data, long long len, float* out, int H, int W, int* src_w, int* src_h);
int idcbh_png_decode_files(const char* const* paths, long long n, float* out, int H, int W, unsigned char* status,
                           int threads);

/* ---- HDF5 subset (Keras `.h5` weight files: load_weights / save_weights, application weight downloads) ----------
 * Reader: superblock v0 / v1 (and v2 / v3 headers), v1 and v2 object headers with continuation blocks, symbol-table
 * groups and compact link messages, contiguous / compact float32 and float64 datasets, fixed- and variable-length
 * string attributes; chunked / compressed datasets and dense link storage are refused with a message. list() writes a
 * JSON description of every dataset (path, shape) and of the string attributes into buf and returns its length
 * (call with buflen 0 to size). */
void* idcbh_h5r_open(const char* path);
long long idcbh_h5r_list(void* h, char* buf, long long buflen);
int idcbh_h5r_read_f32(void* h, const char* path, float* dst, long long n);
void idcbh_h5r_close(void* h);
/* Writer: groups are created on demand from the dataset path; attr_strs writes a fixed-length string array
 * (`scalar` != 0: a single scalar string), the form Keras uses for layer_names / weight_names / backend. */
void* idcbh_h5w_open(const char* path);
int idcbh_h5w_dataset_f32(void* h, const char* path, const float* data, const long long* dims, int rank);
int idcbh_h5w_attr_strs(void* h, const char* group_path, const char* name, const char* const* strs, int n, int scalar);
int idcbh_h5w_close(void* h);          /* writes the file; the handle is gone afterwards, also on failure */

/* ---- checkpoint container (ModelCheckpoint(save_weights_only=True) / load_weights, fed_model.py:186-197) ---------
 * One file: a JSON header (names, shapes, payload offsets) followed by raw little-endian float32 payloads; written
 * to a temporary name and renamed, so a crash never leaves a half-written checkpoint under the final name. */
void* idcbh_ckpt_begin(const char* path);
int idcbh_ckpt_add_f32(void* h, const char* name, const float* data, const long long* shape, int ndim);
int idcbh_ckpt_end(void* h);
long long idcbh_ckpt_header(const char* path, char* buf, long long buflen);   /* header length, or -1 */
int idcbh_ckpt_read(const char* path, unsigned long long header_len, unsigned long long offset, void* dst,
                    unsigned long long nbytes);

/* ---- federated client placement: longest-processing-time-first over `n_ranks` GPUs ----------------------------
 * cost[c] = examples of client c; rank_of_client[c] receives its GPU, load_of_rank (optional) the summed cost. */
int idcbh_schedule_clients(int n_clients, const long long* cost, int n_ranks, int* rank_of_client,
                           long long* load_of_rank);

/* ---- secure aggregation on the host: the oracle of csrc/comm.cu:secagg_kernel (same Philox4x32-10, same ring) ----
 * mask: y[k][i] = Q(w[k][i]) + sum_{j>k} PRG(k,j,i) - sum_{j<k} PRG(j,k,i) (mod 2^32); unmask_mean: the ring sum of
 * the y rows, de-quantised and divided by the client count. n must be a multiple of 4, frac_bits in 1..30. */
int idcbh_secagg_mask(const float* w, uint32_t* y, int n_clients, long long n, int frac_bits, unsigned long long seed);
int idcbh_secagg_unmask_mean(const uint32_t* y, float* out, int n_clients, long long n, int frac_bits);
void idcbh_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out4);

#ifdef __cplusplus
}
#endif
#endif /* IDCB200_HOST_H */
