/*
 * fm_b200.h -- C ABI of libfm_b200.so: the fast_tffm sparse-FM hot path
 * (parse -> unique -> lookup -> score -> loss -> grad -> sparse Adagrad) as
 * hand-written sm_100a CUDA kernels.
 *
 * Conventions (they hold for every entry point unless it says otherwise):
 *   - every pointer is a DEVICE pointer unless named host_*; the caller
 *     allocates every input, output and workspace; the library never
 *     allocates, frees or synchronises;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*), on the
 *     device that is current on the calling thread;
 *   - the return value is 0 on success or a negative FM_E_* code; a message
 *     for the calling thread's last failure is at fm_last_error();
 *   - entry points hold no state between calls and may be called from several
 *     host threads on different streams;
 *   - "file:line" citations point into the reference tree (douban/fast_tffm).
 */
#ifndef FM_B200_H_
#define FM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FM_B200_ABI_VERSION 1

enum {
  FM_OK = 0,
  FM_E_INVALID = -1, /* bad argument (null pointer, negative size, ...) */
  FM_E_WORKSPACE = -2, /* workspace too small */
  FM_E_CUDA = -3,    /* a CUDA runtime call failed; see fm_last_error() */
  FM_E_LIMIT = -4    /* size beyond what the entry point supports */
};

/* parser error kinds written to fm_parse_status.error_kind; the text each one
 * maps to is the reference's (cc/fm_parser_op.cc:70, :77/:88, :90-92, :100) */
enum {
  FM_PARSE_OK = 0,
  FM_PARSE_E_LABEL = 1,    /* "Label could not be read in example: " */
  FM_PARSE_E_FORMAT = 2,   /* "Invalid format in example: " */
  FM_PARSE_E_RANGE = 3,    /* "Invalid feature id. Should be in range [0, vocabulary_size)." */
  FM_PARSE_E_VALUE = 4,    /* "Invalid feature value. " */
  FM_PARSE_E_CAPACITY = 5  /* not a reference error: feature_ids/vals capacity too small */
};

enum { FM_LOSS_LOGISTIC = 0, FM_LOSS_MSE = 1 };

/* device-resident result of a parse.  first_error packs (line << 3) | kind of
 * the FIRST failing line (the reference keeps the first status too,
 * cc/fm_parser_op.cc:39-45); UINT64_MAX means the batch parsed cleanly. */
typedef struct {
  uint64_t first_error;
  int64_t nnz;      /* total number of features = feature_poses[n_lines] (fm_parse_count) */
  int64_t n_lines;  /* written by fm_split_lines */
  int64_t reserved;
} fm_parse_status;
#define FM_PARSE_NO_ERROR UINT64_MAX
#define FM_PARSE_ERROR_KIND(first_error) ((first_error) == UINT64_MAX ? 0 : (int)((first_error) & 7))
#define FM_PARSE_ERROR_LINE(first_error) ((first_error) == UINT64_MAX ? (int64_t)-1 : (int64_t)((first_error) >> 3))

int fm_abi_version(void);
const char* fm_last_error(void);

/* ---- P: FmParser, cc/fm_parser_op.cc:8-15 (op), :29-51 (Compute), :58-109 (ParseLine) ----
 *
 * Text layout: line i of the batch is text[line_begins[i] .. line_ends[i] + end_adjust).
 * For a buffer that holds the lines back to back, each followed by one terminator byte
 * ('\n' in a file) -- the layout fm_split_lines produces -- pass line_begins = line_offsets,
 * line_ends = line_offsets + 1, end_adjust = -1.  Any other selection or order of lines
 * (a shuffled batch) is two gathered index arrays over the same text; the text itself is
 * never moved.  A NUL inside a line ends it (std::string::c_str(), :65).
 * `text` must be 16-byte aligned and readable up to the next multiple of 16 past
 * text_bytes (the kernels read whole 16-byte words).
 */

/* Finds the line starts of a '\n'-separated buffer (the TextLineReader's job,
 * tffm/fm_model.py:40-43, on the device).  A last line without '\n' is NOT
 * accepted: the caller appends one.  Writes line_offsets[0..n_lines] and
 * status->n_lines; at most max_lines lines (more => FM_PARSE_E_CAPACITY). */
size_t fm_split_lines_workspace(int64_t text_bytes);
int fm_split_lines(const uint8_t* text, int64_t text_bytes, int64_t* line_offsets,
                   int64_t max_lines, fm_parse_status* status, void* workspace,
                   size_t workspace_bytes, void* stream);

/* Pass 1: per line, the label-free feature count -> sizes[n_lines]; then
 * feature_poses = [0, cumsum(sizes)] (this is tf.cumsum at
 * tffm/fm_model.py:73 fused into the parser) and status->nnz.
 * Resets status->first_error first.  Lines the fast counter cannot prove regular are
 * counted by a sequential kernel that follows ParseLine statement by statement. */
size_t fm_parse_workspace(int64_t n_lines);
int fm_parse_count(const uint8_t* text, int64_t text_bytes, const int64_t* line_begins,
                   const int64_t* line_ends, int64_t end_adjust, int64_t n_lines,
                   int64_t vocab_size, int hash_feature_id, int32_t* sizes,
                   int32_t* feature_poses /* [n_lines+1] */, fm_parse_status* status,
                   void* workspace, size_t workspace_bytes, void* stream);

/* Pass 2: labels[n_lines], feature_ids[nnz] (int64, as the op's output),
 * feature_vals[nnz]; `capacity` = elements available in feature_ids/vals.
 * `workspace` is the one given to fm_parse_count for this batch (its contents are
 * still needed: the two passes share their bookkeeping).
 * max_blocks: 0 = one thread block per 32 lines; > 0 caps the grid (the blocks then loop),
 * which turns the pass into a thin background kernel when it runs on a side stream next to
 * the memory-bound table kernels of the previous batch (2 blocks per SM works well). */
int fm_parse_fill(const uint8_t* text, int64_t text_bytes, const int64_t* line_begins,
                  const int64_t* line_ends, int64_t end_adjust, int64_t n_lines,
                  int64_t vocab_size, int hash_feature_id, const int32_t* feature_poses,
                  float* labels, int64_t* feature_ids, float* feature_vals, int64_t capacity,
                  fm_parse_status* status, void* workspace, size_t workspace_bytes, int max_blocks,
                  void* stream);

/* ---- U1: tf.unique, tffm/fm_model.py:72 ----
 * ori_ids[U] = distinct ids in order of first occurrence; local_ids[n] = index
 * of every occurrence into ori_ids; *n_unique (device int64) = U.
 * n may come from the device: if n_dev != NULL the kernels read *n_dev (<= n)
 * instead of n, so the parser's nnz need not visit the host. */
size_t fm_unique_workspace(int64_t n);
int fm_unique(const int64_t* ids, int64_t n, const int64_t* n_dev, int64_t* ori_ids,
              int32_t* local_ids, int64_t* n_unique, void* workspace, size_t workspace_bytes,
              void* stream);

/* ---- L1: tf.nn.embedding_lookup, tffm/fm_model.py:291 ----
 * The table lives in ONE device array [vocab_rows, table_ld] (row id = feature
 * id; the reference's vocab_block_b[id / nb] with b = id % nb, :269-284, is a
 * checkpoint-time relabelling).  out[u, 0..K] = table[ori_ids[u], 0..K].
 * ori_ids are NOT range-checked (neither are the reference op's inputs).  guard
 * (optional, device): the call is a no-op unless *guard == UINT64_MAX; pass
 * &status->first_error so the ids of a batch that failed to parse are never used. */
int fm_lookup(const float* table, int64_t table_ld, const int64_t* ori_ids, int64_t n_unique,
              const int64_t* n_unique_dev, int64_t factor_num, float* out, int64_t out_ld,
              const uint64_t* guard, void* stream);

/* ---- S: FmScorer, cc/fm_scorer_op.cc:9-17 (op), cc/fm_scorer_op.h:8-140 ----
 * pred[i] = sum_j b_j x_j + 0.5 sum_k [ (sum_j v_jk x_j)^2 - sum_j (v_jk x_j)^2 ]
 * reg     = 0.5 lf sum_occ sum_k v^2 + 0.5 lb sum_occ b^2   (per occurrence)
 * params is [rows, params_ld], column 0 the bias, 1..K the factors.
 * workspace: fm_score_workspace(batch) bytes (per-block partials of reg). */
size_t fm_score_workspace(int64_t batch);
int fm_score(const int32_t* feature_ids, const float* params, int64_t params_ld,
             const float* feature_vals, const int32_t* feature_poses, int64_t batch,
             int64_t factor_num, float factor_lambda, float bias_lambda, float* pred_score,
             float* reg_score, void* workspace, size_t workspace_bytes, void* stream);

/* fm_score without a workspace argument.  It has no counterpart in the reference.
 * For hosts that cannot allocate device memory for us -- a serving framework that
 * reserved the whole GPU for its own pool so a plugin's cudaMalloc fails or is
 * forbidden, or a latency-bound scoring thread that must stay clear of
 * cudaMalloc/cudaFree's device-wide lock and implicit synchronisation.  The
 * per-block scratch fm_score needs sits in a fixed device array inside the
 * library (512 KiB, held by every GPU that loads the library until exit), so:
 *   - batch <= FM_SCORE_NOALLOC_MAX_BATCH, else FM_E_LIMIT (use fm_score);
 *   - all calls on a device share that array and it keeps the last call's
 *     contents; calls that overlap on the GPU race -- serialise them;
 *   - the first call on each device resolves the array's device address
 *     (cudaGetSymbolAddress); make it at start-up, before any stream capture.
 * Allocates nothing, frees nothing, never synchronises.  Same results as fm_score. */
#define FM_SCORE_NOALLOC_MAX_BATCH 1048576
int fm_score_noalloc(const int32_t* feature_ids, const float* params, int64_t params_ld,
                     const float* feature_vals, const int32_t* feature_poses, int64_t batch,
                     int64_t factor_num, float factor_lambda, float bias_lambda,
                     float* pred_score, float* reg_score, void* stream);

/* ---- LS: loss and d loss / d pred, tffm/fm_model.py:311-316, :324-326 ----
 * loss = mean_i w_i l(pred_i, y_i); pred_grad_i = w_i l'(pred_i, y_i) / batch.
 * weights may be NULL (all ones, :66).  loss is a device float. */
size_t fm_loss_workspace(int64_t batch);
int fm_loss_grad(const float* pred, const float* labels, const float* weights, int64_t batch,
                 int loss_kind, float* loss, float* pred_grad, void* workspace,
                 size_t workspace_bytes, void* stream);

/* ---- G: FmGrad, cc/fm_grad_op.cc:15-24 (op), cc/fm_grad_op.h:23-163 ----
 * params_grad has the shape of params ([param_rows, params_ld], fm_grad_op.cc:67)
 * and is zero-filled by this call (fm_grad_op.h:155) before the accumulation.
 * reg_grad is the op's scalar input; if reg_grad_dev != NULL the kernel reads
 * the scalar from the device instead (autograd hands it over as a tensor).
 * If param_rows_dev != NULL only the first *param_rows_dev (<= param_rows) rows
 * are in use and need clearing. */
int fm_grad(const int32_t* feature_ids, const float* params, int64_t params_ld,
            const float* feature_vals, const int32_t* feature_poses, int64_t batch,
            int64_t factor_num, float factor_lambda, float bias_lambda, const float* pred_grad,
            float reg_grad, const float* reg_grad_dev, int64_t param_rows,
            const int64_t* param_rows_dev, float* params_grad, void* stream);

/* ---- A1: TF1 SparseApplyAdagrad as emitted for tffm/fm_model.py:341-348 ----
 * for u < U: row = ori_ids[u]:  acc[row] += g[u]^2 ; table[row] -= lr g[u] / sqrt(acc[row])
 * guard (optional, device): the call is a no-op unless *guard == UINT64_MAX.  Pass
 * &status->first_error so that a batch that failed to parse never touches the table
 * (in the reference the failed op aborts the step, cc/fm_parser_op.cc:70). */
int fm_sparse_adagrad(float* table, float* accum, int64_t table_ld, const int64_t* ori_ids,
                      int64_t n_unique, const int64_t* n_unique_dev, int64_t factor_num,
                      const float* grad, int64_t grad_ld, float learning_rate, const uint64_t* guard,
                      void* stream);

/* ---- fused training path: S + LS + G + A1 with every table row read once ----
 * Not an op of the reference: it computes what fm_lookup -> fm_score ->
 * fm_loss_grad -> fm_grad -> fm_sparse_adagrad compute (tffm/fm_model.py:291-348
 * over cc/fm_scorer_op.h and cc/fm_grad_op.h), in two kernels.
 *
 * fm_fused_score_scatter: per example, gathers its rows from `table` by GLOBAL id
 * (feature_ids, int64 as the parser emits them), writes pred_score, the mean
 * weighted loss and reg_score, and adds the example's share of the gradient to
 * gacc[local_ids[j]] (local_ids from fm_unique).  A gacc row holds, in columns
 * 0, 1..K, K+1, K+2: C = sum pg x, B_k = sum pg x s_k, D = sum pg x^2, cnt; from
 * them the row's gradient is G_0 = C + cnt rg lb b, G_k = v_k (cnt rg lf - D) + B_k,
 * which equals the sum of FmGrad's per-occurrence terms (cc/fm_grad_op.h:88-97).
 * gacc must be all zero on entry; fm_fused_adagrad leaves it all zero again.
 * table_ld and gacc_ld (row strides) are multiples of 4 floats, columns K+1 up to the
 * next multiple of 4 of a table row are zero, table / gacc / accum are 16-byte aligned.
 *
 * fm_fused_adagrad: for u < U, forms G from gacc[u] and table[ori_ids[u]] and
 * applies TF1 SparseApplyAdagrad to table / accum; rg = reg_grad (the factor the
 * objective puts on reg_score: 1 / batch_size, tffm/fm_model.py:345-347).
 * guard: as in fm_sparse_adagrad; both calls are no-ops when it is tripped.
 * labels == NULL and gacc == NULL together: scoring only (pred_score and reg_score; no
 * loss, no scatter) -- the predict path, reading rows from the table by feature id.
 * loss_batch: denominator of the mean loss and of d loss / d pred; 0 = batch.  A rank
 * that holds a slice of a global batch passes the global size and sums the per-rank
 * losses afterwards. */
int64_t fm_fused_gacc_ld(int64_t factor_num);
size_t fm_fused_workspace(int64_t batch);
int fm_fused_score_scatter(const int64_t* feature_ids, const int32_t* local_ids,
                           const float* feature_vals, const int32_t* feature_poses,
                           const float* labels, const float* weights, int64_t batch,
                           int64_t factor_num, const float* table, int64_t table_ld,
                           float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                           float* pred_score, float* loss, float* reg_score, float* gacc,
                           int64_t gacc_ld, const uint64_t* guard, void* workspace,
                           size_t workspace_bytes, void* stream);
int fm_fused_adagrad(float* table, float* accum, int64_t table_ld, const int64_t* ori_ids,
                     int64_t n_unique, const int64_t* n_unique_dev, int64_t factor_num,
                     float* gacc, int64_t gacc_ld, float learning_rate, float factor_lambda,
                     float bias_lambda, float reg_grad, const uint64_t* guard, void* stream);

/* The same pair without tf.unique: when local_ids == NULL, fm_fused_score_scatter adds
 * to the accumulator row of the feature id itself (gacc[feature_ids[j]]: a table-shaped
 * accumulator), and fm_fused_adagrad_claim walks the n OCCURRENCES `rows` (= feature_ids):
 * the first lane group to swap a row's cnt column to zero applies the update, later
 * occurrences of that row see cnt == 0 and leave.  gtab[row] belongs to table[row].
 * guard tripped: discard_on_guard == 0 -> the call is a no-op (`rows` may be garbage);
 * != 0 -> `rows` is trusted, the accumulated sums of those rows are thrown away and the
 * table is left alone (a step that some other rank voided). */
int fm_fused_adagrad_claim(float* table, float* accum, int64_t table_ld, const int64_t* rows,
                           int64_t n, const int64_t* n_dev, int64_t factor_num, float* gtab,
                           int64_t gtab_ld, float learning_rate, float factor_lambda,
                           float bias_lambda, float reg_grad, const uint64_t* guard,
                           int discard_on_guard, void* stream);

/* ---- sorted training path: G + A1 as a sort / segment reduce fused with the optimizer step ----
 * Not ops of the reference: together with fm_sorted_score they compute what FmGrad's scatter
 * (cc/fm_grad_op.h:80-106; CAS-loop float adds, :3-16) followed by the embedding_lookup gradient and
 * SparseApplyAdagrad (tffm/fm_model.py:341-348) compute, without atomics and without a dense gradient.
 *
 * fm_sort_by_row: stable LSD radix sort of the batch's occurrences by table row.  Needs only the parser's
 *   outputs, so it can run ahead on a side stream next to the parse.  n = *nnz_dev (<= capacity).
 *   sorted_keys[n]   the row keys ascending (key = feature id; with n_ranks > 1: (id % n_ranks) <<
 *                    row_bits | id / n_ranks, row_bits = bit length of ceil(vocab_size / n_ranks) - 1, i.e.
 *                    grouped by owning rank of an `id % n_ranks` partition, rows ascending inside)
 *   sorted_pairs[2n] (example index, bits of feature_vals) of each occurrence, in that order; occurrences
 *                    of one row keep their batch order (the sort is stable => sums are reproducible)
 *   counts[2]        device int64: counts[0] = n (written here), counts[1] = number of distinct keys (= tf.unique's
 *                    U; written by fm_sorted_update, which counts the runs it walks)
 *   guard tripped (*guard != UINT64_MAX): counts = 0, nothing sorted.
 *   vocab_size <= 2^32, capacity < 2^30.  max_blocks: grid cap (0 = default), as in fm_parse_fill.
 * fm_sorted_score: fm_fused_score_scatter's scorer + loss; instead of scattering it leaves one record per
 *   example, records[i] = [pg_i, s_i1 .. s_iK] (pg = d loss / d pred_i / loss_batch; s_ik = sum_j x_j v_jk,
 *   cc/fm_grad_op.h:42-47), record_ld >= fm_sorted_record_ld(K) floats.
 * fm_sorted_update: one lane group per run of equal keys (runs longer than FM_SORT_CHUNK occurrences are cut into
 *   pieces whose partial sums are added up, in order, by a second short kernel -- hot rows under skewed ids):
 *   G_0 = C + cnt rg lb b, G_k = v_k (cnt rg lf - D) + B_k from the
 *   run's occurrences (C = sum pg x, D = sum pg x^2, B_k = sum pg x s_k), then the optimizer step on
 *   table / accum rows (key & (2^row_bits - 1); row_bits = 0: the key is the row).  FM_OPT_ADAGRAD = TF1
 *   SparseApplyAdagrad; FM_OPT_SGD (w -= lr G; accum unused, may be NULL) is an extension -- the reference only
 *   ever builds AdagradOptimizer (tffm/fm_model.py:341-344).  guard as in fm_sparse_adagrad. */
#define FM_SORT_CHUNK 32
enum { FM_OPT_ADAGRAD = 0, FM_OPT_SGD = 1 };
size_t fm_sort_workspace(int64_t capacity);
int fm_sort_by_row(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                   int64_t batch, int64_t capacity, const int64_t* nnz_dev, int64_t vocab_size,
                   int64_t n_ranks, uint32_t* sorted_keys, uint32_t* sorted_pairs,
                   int64_t* counts, const uint64_t* guard, void* workspace, size_t workspace_bytes,
                   int max_blocks, void* stream);
int64_t fm_sorted_record_ld(int64_t factor_num);
int fm_sorted_score(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                    const float* labels, const float* weights, int64_t batch, int64_t factor_num,
                    const float* table, int64_t table_ld, float factor_lambda, float bias_lambda,
                    int loss_kind, int64_t loss_batch, float* pred_score, float* loss, float* reg_score,
                    float* records, int64_t record_ld, const uint64_t* guard, void* workspace,
                    size_t workspace_bytes, void* stream);
size_t fm_sorted_workspace(int64_t capacity, int64_t factor_num);
int fm_sorted_update(float* table, float* accum, int64_t table_ld, const uint32_t* sorted_keys,
                     const uint32_t* sorted_pairs, int64_t* counts, int64_t capacity, const float* records, int64_t record_ld, int64_t factor_num,
                     int optimizer, float learning_rate, float factor_lambda, float bias_lambda,
                     float reg_grad, int64_t row_bits, const uint64_t* guard, void* workspace,
                     size_t workspace_bytes, void* stream);

/* ---- D1: the parameter-server sharding of tffm/fm_model.py:269-291 + run_tffm.py:181-206,
 * as explicit routing around an all-to-all: rank id % n_ranks owns table row id / n_ranks ----
 * fm_shard_bucket: stable counting sort of `ids` (the batch's distinct ids) by owner:
 *   send_ids[n] grouped by owner, pos[u] = position of ids[u] in send_ids, counts[n_ranks].
 * fm_shard_remap: slot64[j] = slot32[j] = pos[local_ids[j]] (row index of occurrence j in
 *   the rows received back / the accumulator rows sent out).
 * fm_shard_local_rows: rows[j] = ids[j] / n_ranks (owner side).
 * fm_shard_combine: dst_rows[index[i], 0:width] += src_rows[i, 0:width] (owner side:
 *   accumulator rows from several ranks for one id are summed before the Adagrad square);
 *   width and both strides are multiples of 4 floats. */
size_t fm_shard_bucket_workspace(int64_t n, int64_t n_ranks);
int fm_shard_bucket(const int64_t* ids, int64_t n, const int64_t* n_dev, int64_t n_ranks,
                    int64_t* send_ids, int32_t* pos, int64_t* counts, void* workspace,
                    size_t workspace_bytes, void* stream);
int fm_shard_remap(const int32_t* local_ids, int64_t n, const int64_t* n_dev, const int32_t* pos,
                   int64_t* slot64, int32_t* slot32, void* stream);
int fm_shard_local_rows(const int64_t* ids, int64_t n, int64_t n_ranks, int64_t* rows, void* stream);
int fm_shard_combine(const float* src_rows, int64_t src_ld, const int64_t* index, int64_t n,
                     int64_t width, float* dst_rows, int64_t dst_ld, void* stream);

/* ---- D1 over NVLink peer memory: no id / row / gradient exchange at all ----
 * Every rank's table shard, accumulator shard and touched-row bitmap are mapped into
 * every process of the box (CUDA IPC / symmetric memory; the *_bases arrays are HOST arrays
 * of n_ranks device pointers, entry r = rank r's shard as mapped here).  Rank id % n_ranks
 * owns row id / n_ranks.
 * fm_peer_score_scatter = fm_fused_score_scatter (table-shaped accumulators) with the row of
 *   every occurrence read from its owner's HBM (128-bit loads over NVLink where remote), the
 *   gradient sums added to the owner's accumulator row with system-scope vector atomics, and
 *   the row's bit set in the owner's bitmap.  labels == NULL: scoring only.
 * All ranks must have finished this call (a stream-ordered barrier, e.g. a tiny all_reduce)
 *   before any owner runs the update:
 * fm_peer_touched_rows: owner side; turns its bitmap into the list of touched rows
 *   (rows[0..*n_rows), ascending) and clears the bitmap.  Feed the list to
 *   fm_fused_adagrad_claim(table_shard, accum_shard, ..., rows, capacity, n_rows, ..., gtab_shard).
 * A second barrier (the loss all_reduce serves) must separate the update from the next
 *   step's reads. */
int fm_peer_score_scatter(const int64_t* feature_ids, const float* feature_vals,
                          const int32_t* feature_poses, const float* labels, const float* weights,
                          int64_t batch, int64_t factor_num, int64_t n_ranks,
                          const void* const* table_bases, int64_t table_ld,
                          void* const* gtab_bases, int64_t gtab_ld, void* const* touched_bases,
                          float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                          float* pred_score, float* loss, float* reg_score, const uint64_t* guard,
                          void* workspace, size_t workspace_bytes, void* stream);
/* fm_peer_sync: the two synchronisation points of that step without a collective library: a barrier over the ranks
 *   of the box that also reduces three values.  Every rank reserves fm_peer_sync_bytes() (32-byte aligned, zeroed
 *   once) inside its peer-mapped allocation; sync_bases is the HOST array of those regions as mapped here.  The call
 *   enqueues one small kernel that publishes {*word_in (+ rank << 48 when it is >= 0: a parser status word),
 *   *a_in, *b_in} to every rank, waits (spinning on system-scope flags) until all n_ranks ranks have published for
 *   this epoch, and writes max(word) to *word_out and the sums of a and b, added in rank order, to sums_out[0..1].
 *   epoch: 1, 2, 3, ... the same sequence on every rank, one per call.  Everything this stream did before the call
 *   (remote atomics included) is visible to every rank's work after its own call returns on the device.
 *   NULL inputs count as -1 / 0.  Replaces the all_reduce pair (tffm has no counterpart: its parameter server
 *   is asynchronous, run_tffm.py:181-206). */
size_t fm_peer_sync_bytes(void);
int fm_peer_sync(void* const* sync_bases, int64_t rank, int64_t n_ranks, uint64_t epoch,
                 const int64_t* word_in, const float* a_in, const float* b_in, int64_t* word_out,
                 float* sums_out, void* stream);
size_t fm_peer_touched_workspace(int64_t shard_rows);

/* ---- D1, record exchange: the same sharded step with the gradients sent as sorted records, not as remote atomics ----
 * (fm_peer_score_scatter puts 192 bytes on NVLink per remote occurrence for 84 bytes of gradient sums.)  A rank sorts
 * its occurrences with fm_sort_by_row(n_ranks = G) -- grouped by owner, rows ascending --, scores with
 * fm_peer_sorted_score (rows pulled from their owners; records [pg, s_1..s_K] per example stay local), then
 * fm_peer_push_records writes, into every owner's inbox (peer-mapped; inbox_bases[r] = rank r's inbox as mapped here):
 *   - per sorted position one key word: the shard row | 2^31 where a piece (a run of equal rows, cut at multiples of
 *     FM_SORT_CHUNK) starts | 2^30 when the run goes on in a further piece (hence at most 2^30 rows per shard);
 *   - per piece one record {C, B_1..B_K, D, cnt} (the columns of fm_fused_score_scatter's accumulator rows), at the
 *     piece's position;
 *   - per tile of fm_merge_tile_rows(G) shard rows the position where its rows begin in this sender's segment.
 * An inbox holds fm_merge_inbox_bytes(inbox_capacity, shard_rows, K, G) bytes (256-byte aligned), one part per sender,
 * each for up to inbox_capacity positions; a segment that does not fit voids the step (status->first_error gets kind
 * FM_PARSE_E_CAPACITY).  All ranks must have finished pushing (fm_peer_sync) before any owner calls
 * fm_peer_merge_update: per tile of rows a block finds every row's records through a directory in shared memory,
 * adds them in registers in sender order (the same sum on every run) and applies TF1 SparseApplyAdagrad to the rows
 * that received something.  flag (device int64, fm_peer_sync's
 * word_out): the call is a no-op when *flag >= 0.  A second fm_peer_sync separates it from the next step's reads.
 * No accumulator table, no bitmap, no global float atomics; about 236 instead of 328 wire bytes per remote occurrence. */
int64_t fm_merge_tile_rows(int64_t n_ranks);
size_t fm_merge_inbox_bytes(int64_t capacity, int64_t shard_rows, int64_t factor_num, int64_t n_ranks);
size_t fm_merge_push_workspace(int64_t n_ranks);
int fm_peer_sorted_score(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                         const float* labels, const float* weights, int64_t batch, int64_t factor_num,
                         int64_t n_ranks, const void* const* table_bases, int64_t table_ld,
                         float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                         float* pred_score, float* loss, float* reg_score, float* records, int64_t record_ld,
                         const uint64_t* guard, void* workspace, size_t workspace_bytes, void* stream);
int fm_peer_push_records(const uint32_t* sorted_keys, const uint32_t* sorted_pairs, const int64_t* counts,
                         int64_t capacity, const float* records, int64_t record_ld, int64_t factor_num,
                         int64_t rank, int64_t n_ranks, int64_t vocab_size, int64_t shard_rows,
                         void* const* inbox_bases, int64_t inbox_capacity, fm_parse_status* status,
                         void* workspace, size_t workspace_bytes, void* stream);
int fm_peer_merge_update(float* table, float* accum, int64_t table_ld, int64_t shard_rows, const void* inbox,
                         int64_t inbox_capacity, int64_t factor_num, int64_t n_ranks, float learning_rate,
                         float factor_lambda, float bias_lambda, float reg_grad, const int64_t* flag,
                         void* stream);
int fm_peer_touched_rows(uint32_t* touched, int64_t shard_rows, int64_t* rows, int64_t capacity,
                         int64_t* n_rows, void* workspace, size_t workspace_bytes, void* stream);
/* fm_peer_touched_rows + fm_fused_adagrad_claim in one kernel, and what the training step uses: the update walks the
 * bitmap itself (a warp per 1024 rows: clears the words, lists the set bits in shared memory, its lane groups take the
 * rows), so every touched row is visited once, in ascending order, with no claim atomics, no scan and no row list.
 * touched: the bitmap fm_peer_score_scatter filled (n_rows bits; any number of ranks, one included: the single-GPU
 * training step is fm_peer_score_scatter with one rank + this call).  guard / discard_on_guard as in
 * fm_fused_adagrad_claim.  rows_seen (optional, device): the number of rows visited is ADDED to it (zero it to get
 * the step's count of distinct rows).  factor_num <= 378 (wider rows: the two-call form above). */
int fm_fused_adagrad_bitmap(float* table, float* accum, int64_t table_ld, uint32_t* touched, int64_t n_rows,
                            int64_t factor_num, float* gtab, int64_t gtab_ld, float learning_rate,
                            float factor_lambda, float bias_lambda, float reg_grad, const uint64_t* guard,
                            int discard_on_guard, int64_t* rows_seen, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FM_B200_H_ */
