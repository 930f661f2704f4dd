/*
 * b2r.h — C ABI of the B200 flat (exact, brute-force) dense-retrieval engine.
 *
 * This is the drop-in boundary for the retrieval step that the reference
 * (Shrinjita/RAG-TL-DomainLLM-Optimizer) consumes but does not implement: the
 * `retrieved_docs` seam at reinforcement_learning_optimization_after_rag.py:286-288
 * (training) and :405 (evaluation), filled in the reference's README tech stack
 * (README.md:27-28) by a FAISS flat index behind LangChain / LlamaIndex.  The
 * reference has no FFI of its own; the entry points below
 * therefore mirror what a FAISS `IndexFlatIP` / `IndexFlatL2` binding would call
 * (`add`, `search(x,k)->(D,I)`, `reset`, `reconstruct`, `normalize_L2`), which is
 * what those retriever stacks bind.  Each declaration names the interface it
 * replaces.
 *
 * Conventions
 *   - plain C, plain pointers and sizes; no C++/torch types cross this boundary;
 *   - every call returns B2R_OK (0) or a negative b2r_status; the text of the last
 *     failure on the calling thread is returned by b2r_last_error();
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *   - data pointers may be HOST or DEVICE pointers (detected with
 *     cudaPointerGetAttributes).  With host pointers the call stages through the
 *     handle's own device buffers and returns after the results are in host memory;
 *     with device pointers the call only enqueues work on `stream`;
 *   - there is no CPU fallback: without a CUDA device every compute entry point
 *     fails with B2R_ERR_CUDA.
 */
#ifndef B2R_H_
#define B2R_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2r_index b2r_index;

typedef enum b2r_status {
  B2R_OK = 0,
  B2R_ERR_ARG = -1,     /* bad argument (shape, k, null pointer, unsupported size) */
  B2R_ERR_CUDA = -2,    /* CUDA runtime / driver failure, or no device             */
  B2R_ERR_OOM = -3,     /* device allocation failed                                 */
  B2R_ERR_STATE = -4,   /* call not valid in the handle's current state             */
  B2R_ERR_LIMIT = -5    /* size beyond a documented fixed limit                     */
} b2r_status;

/* FAISS MetricType values (faiss METRIC_INNER_PRODUCT = 0, METRIC_L2 = 1). */
enum { B2R_METRIC_IP = 0, B2R_METRIC_L2 = 1 };

/* Corpus storage / arithmetic mode.
 *   B2R_STORE_F32 : "exact" mode.  fp32 rows are kept; candidates come from an
 *                   fp32-accurate scoring pass and are re-scored with the fixed fp64
 *                   arithmetic documented in DESIGN.md §3, so ids and scores are a
 *                   pure function of the inputs (ties: lower id first).
 *   B2R_STORE_BF16: "fast" mode.  rows and queries are rounded to bf16 (RN-even);
 *                   the result is the exact search over those rounded vectors. */
enum { B2R_STORE_F32 = 0, B2R_STORE_BF16 = 1 };

/* Scoring engine selection for b2r_index_set_engine (0 is the default). */
enum {
  B2R_ENGINE_AUTO = 0,  /* tcgen05 tensor-core kernel when the shape allows, else FMA */
  B2R_ENGINE_FMA = 1,   /* CUDA-core fp32 FMA streaming kernel                        */
  B2R_ENGINE_UMMA = 2,  /* TMA + tcgen05.mma + TMEM kernel: kind::f16 on bf16 rows,   */
                        /* one kind::tf32 pass on fp32 rows                           */
  B2R_ENGINE_UMMA3 = 3, /* the same kernel with the 3xTF32 split (fp32 rows); also    */
                        /* what exact mode's second (range) pass runs                 */
  B2R_ENGINE_SHADOW = 4 /* reported only: kind::f16 on the bf16 shadow of fp32 rows   */
                        /* (b2r_index_set_shadow), exact re-scoring on the fp32 rows  */
};

/* Counters of the most recent search on a handle (b2r_index_last_stats). */
typedef struct b2r_search_stats {
  int64_t nq;             /* queries in the call                                        */
  int64_t k;              /* requested k                                                */
  int64_t kprime;         /* candidates kept per query before exact re-scoring          */
  int64_t unverified;     /* queries whose top-k is NOT proven complete when the call   */
                          /* is over.  Always 0 by default: a query the first pass      */
                          /* cannot prove is re-searched on the device */
                          /* (range pass, then the exhaustive scan, which cannot fail). */
                          /* Non-zero only after b2r_index_set_verify(idx, 0).          */
  int64_t candidates;     /* candidate entries that reached the merge kernels           */
  int32_t engine;         /* B2R_ENGINE_FMA / _UMMA / _UMMA3 / _SHADOW of the first pass */
  int32_t launches;       /* kernels launched by the call                               */
  int32_t fallback;       /* 1 if any query needed the second pass                      */
  int32_t reserved;
  int64_t main_rows;      /* rows covered by the launch(es) that b2r_index_score_time times: */
                          /* the main tcgen05 launch (all rows but the seed rows), or all    */
                          /* rows for the FMA engine                                         */
  int64_t fallback_queries;   /* queries the first pass could not prove: only these are     */
                              /* searched again (range pass: every row above a fixed bound)  */
  int64_t exhaustive_queries; /* of those, queries that went on to the exhaustive DOT64 scan */
                              /* (their candidate set did not fit: thousands of tied rows)   */
} b2r_search_stats;

/* Library / device ------------------------------------------------------------------ */

/* Text of the last error raised on the calling thread ("" if none).
 * Replaces: faiss c_api `faiss_get_last_error()`. */
const char* b2r_last_error(void);

/* Library version string, e.g. "b2r 0.1 sm_100a". */
const char* b2r_version(void);

/* Number of CUDA devices visible (0 when there is none; never fails). */
int b2r_device_count(void);

/* Index life cycle ------------------------------------------------------------------ */

/* Create an empty flat index of dimension d on CUDA device `device`.
 * Replaces: faiss `IndexFlatIP(d)` / `IndexFlatL2(d)` construction
 * (c_api `faiss_IndexFlat_new_with`), the vector store named at README.md:28. */
int b2r_index_create(b2r_index** out, int64_t d, int metric, int storage, int device);

/* Destroy the index and every device buffer it owns.  Replaces `faiss_Index_free`. */
void b2r_index_free(b2r_index* idx);

/* Append n rows (fp32, row-major, contiguous [n,d]); ids continue from ntotal.
 * If `normalize` is non-zero each row is L2-normalised on the way in (fused with
 * the bf16 cast in fast mode); zero-norm rows are stored unchanged.
 * Replaces: faiss `Index::add(n, x)`; LangChain FAISS.add_embeddings /
 * LlamaIndex FaissVectorStore.add call exactly this. */
int b2r_index_add(b2r_index* idx, int64_t n, const float* x, int normalize, void* stream);

/* Reserve device storage for `n` rows up front (optional; add() grows geometrically). */
int b2r_index_reserve(b2r_index* idx, int64_t n);

/* Remove all rows, keep the handle.  Replaces faiss `Index::reset()`. */
int b2r_index_reset(b2r_index* idx);

/* Number of stored rows.  Replaces faiss `Index::ntotal`. */
int b2r_index_ntotal(const b2r_index* idx, int64_t* out);

/* Copy stored row i back as fp32 (the stored value: normalised / bf16-rounded).
 * Replaces faiss `Index::reconstruct(i, out)`. */
int b2r_index_reconstruct(b2r_index* idx, int64_t i, float* out, void* stream);

/* Copy n stored rows starting at row0 out in their STORED representation: dense [n,d] float32
 * (exact mode) or bfloat16 (fast mode), exactly the values search() scores.  `out` may be a host or
 * a device pointer.  With b2r_index_import_rows this is the per-rank shard save/load
 * (SURVEY.md §8 f3); the faiss counterpart is `write_index` / `read_index` on an IndexFlat. */
int b2r_index_export_rows(b2r_index* idx, int64_t row0, int64_t n, void* out, void* stream);

/* Append n rows given in the STORED representation (as b2r_index_export_rows wrote them): no
 * normalisation, no rounding; row norms are recomputed on the device. */
int b2r_index_import_rows(b2r_index* idx, int64_t n, const void* rows, void* stream);

/* Search ---------------------------------------------------------------------------- */

/* Exact top-k over all stored rows for nq queries (fp32 [nq,d]).
 * D: float32 [nq,k], I: int64 [nq,k], best first: inner product descending, or
 * SQUARED L2 distance ascending; equal score -> lower id first.  When k > ntotal
 * the tail is padded with I=-1 and D=-FLT_MAX (IP) / +FLT_MAX (L2).
 * `row_offset` is added to every returned id (global id of a row-sharded corpus).
 * If `normalize_queries` is non-zero a normalised copy of the queries is searched.
 * Calls with more than 8192 queries are walked in batches of 8192 (bounded scratch).
 * Replaces: faiss `Index::search(n, x, k, distances, labels)` — the call made by
 * LangChain `FAISS.similarity_search_with_score_by_vector` and LlamaIndex
 * `FaissVectorStore.query`, whose output feeds
 * reinforcement_learning_optimization_after_rag.py:287 / :31 / :63. */
int b2r_index_search(b2r_index* idx, int64_t nq, const float* q, int64_t k,
                     float* D, int64_t* I, void* stream);

int b2r_index_search_ex(b2r_index* idx, int64_t nq, const float* q, int64_t k,
                        int normalize_queries, int64_t row_offset,
                        float* D, int64_t* I, void* stream);

/* Shard-local search whose result is left PACKED for a cross-GPU merge: one
 * uint64 per (query, rank) = order-preserving score bits << 32 | ~(uint32 global id),
 * larger = better; empty slots are 0.  packed: uint64 [nq,k] (device pointer).
 * This is the per-rank payload of the single all-gather. */
int b2r_index_search_packed(b2r_index* idx, int64_t nq, const float* q, int64_t k,
                            int normalize_queries, int64_t row_offset,
                            uint64_t* packed, void* stream);

/* b2r_merge_packed for lists that are g_stride (>= nq*k) elements apart (payloads that carry
 * extra words behind the nq*k hits).  At most 1024 lists. */
int b2r_merge_packed_strided(int metric, int64_t G, int64_t nq, int64_t k, const uint64_t* packed,
                             int64_t g_stride, float* D, int64_t* I, int device, void* stream);

/* Merge G gathered packed lists ([G,nq,k], device) into the final D/I ([nq,k]).
 * The comparator is the one of b2r_index_search, so a row-sharded search equals the
 * unsharded one.  No counterpart in faiss CPU; it is what `IndexShards` does on host. */
int b2r_merge_packed(int metric, int64_t G, int64_t nq, int64_t k,
                     const uint64_t* packed, float* D, int64_t* I, int device, void* stream);

/* Row-sharded search: one process (or thread) and one handle per GPU, rank r holding the
 * rows [row_offset, row_offset + ntotal) of the global corpus; every rank calls
 * b2r_index_search_sharded with the SAME queries and k (collective semantics).  Each call enqueues, on
 * `stream`: the shard-local search (packed top-k with global ids), ONE ncclAllGather of the
 * nq*k*8-byte payloads into a buffer the handle owns and keeps, and the merge kernel — the result
 * equals the unsharded search on every rank, ties included.  No host synchronisation for device
 * pointers.  No faiss-CPU counterpart (IndexShards does the same merge on the host).
 *
 * b2r_index_set_comm adopts a communicator the caller created (an ncclComm_t, passed as void*; it must
 * come from the NCCL library this process has loaded, which is the one b2r binds with dlopen); the
 * handle does not destroy it.  b2r_index_comm_init creates one from a 128-byte ncclUniqueId that
 * rank 0 obtained with b2r_nccl_unique_id and sent to the other ranks by any means; the handle owns
 * and destroys it.  world == 1 needs no communicator (the call degenerates to a local search). */
int b2r_nccl_unique_id(void* out128);
int b2r_index_set_comm(b2r_index* idx, void* nccl_comm, int rank, int world, int64_t row_offset);
int b2r_index_comm_init(b2r_index* idx, const void* unique_id128, int rank, int world, int64_t row_offset);
int b2r_index_search_sharded(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries,
                             float* D, int64_t* I, void* stream);

/* Tuning / introspection ------------------------------------------------------------ */

/* Force a scoring engine (B2R_ENGINE_*); AUTO picks per call. */
int b2r_index_set_engine(b2r_index* idx, int engine);

/* Exact mode only, before the first add: also keep a bf16 copy of the rows (+50 % memory) and let the
 * scoring pass run on it — half the HBM bytes and twice the MMA rate of the TF32 pass; results stay
 * those of exact mode (fp32 rows are re-scored, the proof and its fallback apply).  No faiss counterpart. */
int b2r_index_set_shadow(b2r_index* idx, int on);

/* By default every search ends with three launches that re-search, on the device and without any
 * host round trip, exactly the queries whose top-k the first pass could not prove (range pass,
 * its merge, exhaustive scan); they return at once when there is none.  on = 0 drops them: unproven
 * queries are then only counted (b2r_index_last_stats: `unverified`) and handling them is the
 * caller's call.  No faiss counterpart (faiss-CPU scores every row in fp32 and has nothing to prove). */
int b2r_index_set_verify(b2r_index* idx, int on);

/* Extra candidates kept per query before the exact re-scoring (default 6). */
int b2r_index_set_pad(b2r_index* idx, int pad);

/* Counters of the last search.  Waits for that search's kernels only (an event recorded behind them),
 * not for the device. */
int b2r_index_last_stats(b2r_index* idx, b2r_search_stats* out);

/* Milliseconds spent in the scoring kernel(s) alone during the last search, measured
 * with CUDA events on the search's stream (synchronises).  Used by bench.py for the
 * roofline figure. */
int b2r_index_last_score_ms(b2r_index* idx, float* ms);

/* Sum of the dominant scoring launches' times (main tcgen05 launch, or all FMA passes; ms,
 * CUDA events on the search's stream; rows covered: b2r_search_stats.main_rows) of the searches
 * made since the last reset (at most the 256 most recent), and how many they were;
 * reset != 0 clears the record afterwards.  Synchronises.  No faiss counterpart: it exists so
 * that bench.py can report the dominant kernel's duration over its own timed region. */
int b2r_index_score_time(b2r_index* idx, int reset, double* total_ms, int64_t* nsearch);

/* Stand-alone kernels ---------------------------------------------------------------- */

/* In-place row-wise L2 normalisation of fp32 [n,d]; zero-norm rows are left unchanged.
 * Replaces: faiss `normalize_L2(x)` (fvec_renorm_L2) and the `normalize(X)` inside
 * sklearn `cosine_similarity`, which the reference calls at
 * reinforcement_learning_optimization_after_rag.py:70, :78 and :104. */
int b2r_normalize_l2(int64_t n, int64_t d, float* x, int device, void* stream);

/* Batched reward cosine: for each of n samples, the cosine of
 * resp[i,:] with each of m rows docs[i,j,:] and the maximum over j
 * (reinforcement_learning_optimization_after_rag.py:63-71).  cos: [n,m] (may be NULL),
 * maxcos: [n] (0.0 when m == 0, as at :71). */
int b2r_cosine_rows(int64_t n, int64_t m, int64_t d, const float* resp, const float* docs,
                    float* cos, float* maxcos, int device, void* stream);

/* Largest n b2r_cosine_rows_noalloc accepts (entries of its fixed device scratch). */
#define B2R_NOALLOC_MAX_SAMPLES 65536

/* b2r_cosine_rows for callers that must not allocate.  No counterpart in the reference.
 * Meant for the reward step inside a training or serving loop whose framework owns all free device
 * memory (a raw cudaMalloc there can fail, and cudaMalloc/cudaFree stall every stream of the device),
 * and for loops that forbid allocation after warm-up.  Same arithmetic and outputs as b2r_cosine_rows,
 * but: DEVICE pointers only (resp [n,d], docs [n,m,d], cos [n,m] or NULL, maxcos [n]); the work is only
 * enqueued on `stream`; n <= B2R_NOALLOC_MAX_SAMPLES, larger n is refused with B2R_ERR_LIMIT.
 * Its scratch is a fixed 256 KiB array that lives on every GPU that has loaded the library until the
 * process exits; the call performs no cudaMalloc / cudaFree or other allocating routine.
 * The first call on each device looks up that array's device address; make that first call outside
 * any stream capture.  All calls on a device share the array (each call clears the part it uses), so
 * two calls whose GPU work overlaps — different streams, or different threads — race: serialise them. */
int b2r_cosine_rows_noalloc(int64_t n, int64_t m, int64_t d, const float* resp, const float* docs,
                            float* cos, float* maxcos, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2R_H_ */
