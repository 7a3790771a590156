// internal.h -- structures shared by the translation units of libphoton_ml_b200.so.
// Nothing here is part of the ABI (include/photon_ml_b200.h is).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/photon_ml_b200.h"

namespace pml {

void set_error(const std::string &msg);
int fail(int code, const std::string &msg);

#define PML_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess)                                                               \
            return ::pml::fail(_e == cudaErrorMemoryAllocation ? PML_ENOMEM : PML_ECUDA,     \
                               std::string(#expr) + ": " + cudaGetErrorString(_e));          \
    } while (0)

#define PML_TRY(expr)               \
    do {                            \
        int _rc = (expr);           \
        if (_rc != PML_OK) return _rc; \
    } while (0)

// ---- geometry of the entry-chunked kernels: a chunk is SEG_CHUNK consecutive entries, owned by one warp ----
constexpr int SEG_CHUNK = 256;               // 32 lanes x 8 entries
#ifndef PML_SEG_TABLE_KB
#define PML_SEG_TABLE_KB 96
#endif
constexpr int SEG_TABLE_BYTES = PML_SEG_TABLE_KB * 1024;   // hot prefix of the coefficient vector staged per CTA (CSR stream)

// ---- geometry of the lanes-per-row / segment-per-warp row kernel ----
constexpr int R_THREADS = 256;
constexpr int R_WARPS = R_THREADS / 32;
constexpr int R_STREAM_CAP = 512;  // products staged per warp per piece in the segment-per-warp variant
constexpr int SMALL_IO_DIM = 4096;   // host vectors up to this length are staged through pinned memory
constexpr int MAX_PARTIALS = 4096; // upper bound on row-kernel CTAs (block partials of F and P)

enum RowMode {
    MODE_VG = 0,        // r_i = wt*l'(z_i);  F += wt*l, P += wt*l'
    MODE_HV_COLD = 1,   // gathers (ec, ev) pairs; r_i = wt*l''(z_i)*(x_i.ev - fs); caches wt*l''
    MODE_HV_CACHED = 2, // gathers ev; r_i = D_i*(x_i.ev - fs)
    MODE_DIAG = 3,      // r_i = wt*l''(z_i), raw coefficients
    MODE_SCORE = 4      // r_i = x_i.c
};

struct RowPassParams {
    int64_t n;
    const uint32_t *row_ptr;  // n+1
    const int32_t *col;       // nnz
    const void *val;          // nnz (float or double)
    const double *label;      // n
    const double *offset;     // n or null
    const double *weight;     // n or null
    const double *vec;        // d doubles (ONE) or d double2 pairs (TWO)
    const double *shift_partials;  // per-CTA partial dots ec.s (and ev.s interleaved when TWO) or null
    int n_shift_partials;
    double *r;                // n: per-row multiplier consumed by the transpose kernel
    double *dcache;           // n: wt*l''(z) written (HV_COLD) or read (HV_CACHED)
    double *partials;         // 2 * gridDim.x: per-CTA (F, P)
    double *row_loss;         // n or null: wt_i * l_i per row (batched per-entity values)
    const unsigned char *block_active;  // (n+31)/32 or null: 0 = every row of the 32-row block belongs to a masked entity
    int mode;
    int loss;
};

struct SegParams {
    int64_t nnz;
    const int32_t *idx;       // nnz: table index (CSC: row, CSR: internal column id); bit 31 = first entry of a run
    const void *val;          // nnz
    const int32_t *run_id;    // run ordinal -> output slot (CSC: original column id, CSR: row id); null = identity
    const int32_t *chunk_lo;  // n_chunks+1: ordinal of the first run starting at or after each chunk
    const double *table;      // what idx looks up (CSC: r, CSR: effective coefficients or (ec, ev) pairs)
    double *out, *out2;       // one slot per run (zeroed beforehand where runs can be absent); out2 with TWO
    double *carry, *carry2;   // n_chunks: sum of the part of a chunk that precedes its first run start
    int n_chunks;
    int n_hot;                // leading table entries staged in shared memory (set by the launcher)
    const unsigned char *active;  // n_chunks or null: 0 = every entry of the chunk belongs to a masked entity (skipped)
};

struct FixEntry {
    int32_t col;          // column id the heads belong to
    int32_t chunk_begin;  // carry_head[chunk_begin .. chunk_end) are added to out[col]
    int32_t chunk_end;
};

struct TiledSide;

// launchers (kernels.cu); each returns the number of kernels it launched, <0 on launch error
int launch_prepare_vectors(cudaStream_t s, int d, const double *coef, const double *v, const double *factors,
                           const double *shifts, const int32_t *rank, double *out_vec, double *shift_partials,
                           int *n_partials);
int launch_row_pass(cudaStream_t s, int sm_count, int value_bytes, int lanes_per_row, bool two, const RowPassParams &p,
                    int *grid_out);
int launch_seg_spmv(cudaStream_t s, int sm_count, int value_bytes, bool two, bool square, SegParams p, int64_t table_len);
int launch_row_epilogue(cudaStream_t s, int sm_count, const RowPassParams &p, const double *za, const double *zb,
                        int *grid_out);
int launch_fixup(cudaStream_t s, const FixEntry *fix, int n_fix, int n_long, const double *carry_head, double *out,
                 const double *carry_head2, double *out2, int d, const double *partials, int n_partials,
                 bool write_value, bool write_prefactor);
int launch_finalize(cudaStream_t s, int d, const double *raw, const double *factors, const double *shifts,
                    double *out_value, double *out_vec);

int launch_seg_tiled(cudaStream_t s, int sm_count, int value_bytes, bool square, const TiledSide &t, const double *r,
                     int64_t n);
int launch_reduce_partial(cudaStream_t s, int n_tiles, int n_cols, const double *partial, const int32_t *nzcol_id,
                          double *out);
int launch_offsets_from_scores(cudaStream_t s, int64_t n, const double *base, const double *scores, const int64_t *row_map,
                               double *out);
int launch_globalize_cols(cudaStream_t s, int64_t n_entities, const int64_t *ent_row_ptr, const int64_t *coef_ptr,
                          const int32_t *ent_dim, const int64_t *row_ptr, int32_t *col, unsigned long long *bad);
int launch_entity_values(cudaStream_t s, int64_t n_entities, const int64_t *ent_row_ptr, const double *row_loss,
                         double *out);
int launch_entity_of_coef(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, int32_t *ent_of);
int launch_segment_dot(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const double *x, const double *y,
                       double *out);
int launch_segment_axpby(cudaStream_t s, int64_t d, const int32_t *ent_of, const double *a, const double *x,
                         const double *c, const double *y, const uint8_t *active, double *out);
int launch_batch_cg_step(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const uint8_t *run, double l2,
                         double *direction, const double *hd, double *step, double *residual, double *rtr,
                         const double *delta, uint8_t *over);
int launch_batch_lbfgs_direction(cudaStream_t s, int64_t n_entities, int64_t d_total, const int64_t *coef_ptr, const uint8_t *go,
                                 int m, const double *S, const double *Y, const double *rho, const int32_t *head,
                                 const int32_t *len, const double *grad, double *out);
int launch_batch_lbfgs_push(cudaStream_t st, int64_t n_entities, int64_t d, const int32_t *ent_of, const uint8_t *mask, int m,
                            const double *rho_new, const double *s, const double *y, double *S, double *Y, double *rho,
                            int32_t *head, int32_t *len);
int launch_check_intercepts(cudaStream_t s, int value_bytes, int64_t n_entities, const int64_t *ent_row_ptr,
                            const uint32_t *row_ptr, const int32_t *col, const void *val, const int32_t *rank,
                            const int64_t *icpt_pos, unsigned long long *bad);
int launch_scale(cudaStream_t s, int64_t d, const double *x, const double *f, double *out);
int launch_absorb_shift(cudaStream_t s, int64_t n_entities, const int64_t *icpt_pos, const double *dots, double *vec);
int launch_batch_finalize(cudaStream_t s, int64_t d, const int32_t *ent_of, const double *V, const double *factors,
                          const double *shifts, const int64_t *icpt_pos, double *out);
int launch_active_flags(cudaStream_t s, int64_t n_entities, const unsigned char *active, const int64_t *ent_row_ptr,
                        int64_t n, int64_t nnz, const uint32_t *row_ptr, int n_row_chunks, unsigned char *row_chunk_active,
                        unsigned char *row_block_active, int64_t m, const uint32_t *nzcol_ptr, const int32_t *nzcol_id,
                        const int32_t *ent_of_coef, int64_t col_nnz, int n_col_chunks, unsigned char *col_chunk_active);
int launch_mask_entities(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const uint8_t *active, double *values,
                         double *vec);

// pack.cu: builds the CSC copy and its chunk metadata on the device
struct Packed;
struct RowSide;
int build_transpose(cudaStream_t s, int sm_count, int64_t n, int32_t d, int64_t nnz, const uint32_t *row_ptr,
                    int32_t *col, const void *val, int value_bytes, int chunk_entries, bool renumber, Packed *out);
constexpr int TILE_MAX_ROWS = 25000;            // r slice + zero slot <= 200 KB of shared memory (one CTA per SM)
constexpr int64_t TILE_MIN_ROWS_DEFAULT = 262144;  // below this r lives happily in L2: no tiling
constexpr int64_t TILE_PARTIAL_BYTES = 128ll << 20; // cap on the (tile, column) partial matrix
constexpr int64_t SUPER_ROWS_DEFAULT = 6ll << 20;   // rows per super block of the cold columns: 48 MB of r in the L2
constexpr int64_t SUPER_MIN_ROWS_DEFAULT = 6ll << 20;  // up to this many rows r (48 MB) is left in one piece
int build_row_side(cudaStream_t s, int64_t n, int64_t nnz, const uint32_t *row_ptr, int32_t *col, int chunk_entries,
                   RowSide *out);
int bank_order(cudaStream_t s, int64_t n_chunks, int32_t *idx, void *val, int value_bytes, int32_t n_table);
int narrow_check(cudaStream_t s, const double *val, int64_t nnz, bool *all_exact);
int ones_check(cudaStream_t s, const void *val, int value_bytes, int64_t nnz, bool *all_ones);
int convert_values(cudaStream_t s, const void *src, int src_bytes, void *dst, int dst_bytes, int64_t nnz);
int check_columns(cudaStream_t s, const int32_t *col, int64_t nnz, int32_t d, int64_t *bad);
int convert_row_ptr(cudaStream_t s, const int64_t *src, uint32_t *dst, int64_t n_plus_1, int64_t *bad);

// Hot columns of the CSC copy, tiled by row blocks so that a CTA gathers r from a slice staged in shared memory.
struct TiledSide {
    uint16_t *idx16 = nullptr;     // owned: row index LOCAL to the tile (tile_rows = the zero slot used by padding),
    void *val = nullptr;           //   bit 15 on the first entry of each (tile, column) run; val: view into val_alloc
    unsigned char *uniform = nullptr;  // owned, n_chunks: 1 = no run starts in the chunk (it lies inside one run)
    int64_t len = 0;               // entries incl. padding; every tile starts at a multiple of 256
    int32_t *run_id = nullptr;     // run ordinal -> slot tile * n_hot_cols + column in `partial`
    int32_t *chunk_lo = nullptr;   // n_chunks + 1
    int32_t *tile_first_chunk = nullptr;  // n_tiles + 1
    FixEntry *fix = nullptr;
    int n_fix = 0, n_fix_long = 0;
    int n_chunks = 0;
    int64_t n_runs = 0;
    int tile_rows = 0, n_tiles = 0, n_hot_cols = 0;
    double *partial = nullptr;     // n_tiles * n_hot_cols per-(tile, column) sums, reduced by reduce_partial_kernel
    double *carry = nullptr;       // n_chunks
};

struct Packed {
    int32_t *idx_alloc = nullptr;  // row indices of the cold region (all columns when nothing is tiled)
    void *val_alloc = nullptr;     // one allocation holds the values of [hot tiled region | cold region]
    TiledSide hot;
    // cold region (all columns when nothing is tiled): plain CSC ordered by (internal column, row)
    int64_t cold_nnz = 0;
    const int32_t *cold_run_id = nullptr;  // run ordinal -> original column id (= nzcol_id + hot.n_hot_cols), or,
                                           //   with super blocks, -> slot of cold_partial (= cold_slot)
    int n_super = 1;                       // super blocks of rows the cold region is cut into (1 = plain CSC)
    int cold_cols = 0;                     // m - hot.n_hot_cols when n_super > 1
    int32_t *cold_slot = nullptr;          // owned; run ordinal -> super block * cold_cols + (column - K)
    double *cold_partial = nullptr;        // n_super * cold_cols per-(super block, column) sums
    int32_t *csc_row = nullptr;    // view: row index, bit 31 set on the first entry of each column
    void *csc_val = nullptr;
    uint32_t *nzcol_ptr = nullptr; // d+1: entry offset of each internal column (columns >= m are empty)
    int32_t *nzcol_id = nullptr;   // d: internal id -> original column id
    int32_t *rank = nullptr;       // d: original column id -> internal id
    int chunk_entries = SEG_CHUNK;
    int32_t *chunk_col_lo = nullptr;
    FixEntry *fix = nullptr;
    int64_t m = 0;  // non-empty columns
    int n_chunks = 0;
    int n_fix = 0;
    int n_fix_long = 0;  // leading fix entries that span >= 32 chunks (one warp each)
    int64_t bytes = 0;
};

// chunk metadata of the CSR stream (the column array itself carries the row-start flags)
// Sliced-ELL copy of the CSR rows (pack.cu build_sell; kernels.cu sell_rows_kernel)
struct SellSide {
    int32_t *idx = nullptr;   // column id + 1, 0 = padding; [slice][quad of positions][row of the slice][4]
    void *val = nullptr;
    int64_t *base = nullptr;  // n_slices + 1: first entry of every slice (multiples of 128)
    int32_t *perm = nullptr;  // n: the row at every position (rows are sorted by length inside windows)
    int64_t n_slices = 0, len = 0, bytes = 0;
};
int build_sell(cudaStream_t s, int64_t n, int64_t nnz, const uint32_t *row_ptr, const int32_t *col, const void *val,
               int value_bytes, double max_overhead, int sigma_shift, int32_t n_hot, SellSide *out);
int launch_sell_rows(cudaStream_t s, int sm_count, int value_bytes, const SellSide &sell, int64_t n, const double *vec,
                     int64_t table_len, double *out, double *out2);

struct RowSide {
    int32_t *nzrow_id = nullptr;  // non-empty row ordinal -> row id; null when no row is empty
    int32_t *chunk_lo = nullptr;
    FixEntry *fix = nullptr;
    int64_t n_runs = 0;
    int n_chunks = 0;
    int n_fix = 0;
    int n_fix_long = 0;  // leading fix entries that span >= 32 chunks (one warp each)
    int64_t bytes = 0;
};

}  // namespace pml

// One rank's end of the peer-memory merge (peer.cu)
struct pml_peer {
    pml_ctx *ctx = nullptr;
    int rank = 0, world = 1;
    int64_t d = 0;
    void *slab = nullptr;                      // [header | raw (d+2) | result (d+1)], cudaMalloc'ed, exported by IPC
    void *peer_slab[PML_MAX_PEERS] = {};       // every rank's slab as mapped here ([rank] = slab)
    bool opened[PML_MAX_PEERS] = {};
    bool connected = false;
    int *h_status = nullptr, *d_status = nullptr;  // mapped host word: 1 = a barrier timed out
    // NVSwitch multicast transport (optional): the slab again, in memory bound to a multicast object
    unsigned long long mc_handle = 0, mc_mem = 0;  // CUmemGenericAllocationHandle
    size_t mc_bytes = 0;
    bool mc_bound = false;
    void *uc_va = nullptr, *mc_va = nullptr;       // this rank's copy / all ranks' copies
};

namespace pml {
int launch_peer_merge(cudaStream_t s, int sm_count, const pml_peer *peer, const double *factors, const double *shifts);
double *peer_raw(const pml_peer *peer);
double *peer_result(const pml_peer *peer);
double *peer_vec(const pml_peer *peer, int which);
void peer_slice(const pml_peer *peer, int64_t *lo, int64_t *hi);
int launch_peer_gather(cudaStream_t s, int sm_count, const pml_peer *peer, int n_vec);
size_t peer_slab_bytes(int64_t d, int world);
}  // namespace pml

struct pml_batch {
    pml_dataset *ds = nullptr;      // the block-diagonal data set (owned)
    int64_t n_entities = 0;
    int64_t *ent_row_ptr = nullptr; // device, E+1
    double *ent_values = nullptr;   // device, E
    std::vector<int64_t> coef_ptr;  // host, E+1
    int64_t *coef_ptr_dev = nullptr; // device, E+1
    int32_t *ent_of_coef = nullptr;  // device, d_total, lazily: entity of every coefficient (segment_axpby)
    uint8_t *active_dev = nullptr;  // device, E; null = every entity active
    // per-entity NormalizationContext (pml_batch_set_normalization); all null = none
    double *factors = nullptr, *shifts = nullptr;  // device, d_total (concatenated)
    int64_t *icpt_pos = nullptr;    // device, E: concatenated index of entity e's intercept coefficient, -1 = none
    double *ec = nullptr, *ev = nullptr;  // device, d_total: effective coefficients / multiply vector of the running call
    double *dots = nullptr;         // device, E
};

struct pml_span {
    int kind;
    cudaEvent_t a, b;
};

struct pml_ctx {
    bool profiling = false;
    std::vector<pml_span> spans;
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    int64_t launches = 0;
    // pinned staging for scalars, and for small coefficient / gradient vectors of host-buffer calls
    double *h_scalar = nullptr;
    double *h_stage = nullptr;  // 2 * (SMALL_IO_DIM + 2) doubles
    std::vector<int *> peer_status;  // mapped host words of this context's peer handles: 1 = an exchange timed out
};

// a captured value+gradient sequence and what it was captured for
// What a captured launch sequence depends on besides the kernels' own inputs: replayed only while all of it matches.
struct pml_graph_key {
    uint64_t version = 0;       // state of the data set (rows, normalization, variant)
    int loss = -1;
    const double *coef = nullptr, *v = nullptr;
    double *out = nullptr;
    int flags = 0;              // sequence-specific switches (want_vector, want_row_loss, cache the curvature ...)
    bool operator==(const pml_graph_key &o) const
    {
        return version == o.version && loss == o.loss && coef == o.coef && v == o.v && out == o.out && flags == o.flags;
    }
};

struct pml_graph {
    cudaGraphExec_t exec = nullptr;
    pml_graph_key key;   // what `exec` was captured for
    pml_graph_key last;  // the most recent request
    int kernels = 0;
    int misses = 0;  // consecutive re-captures caused by nothing but different buffers
};

struct pml_dataset {
    pml_ctx *ctx = nullptr;
    int64_t n = 0, nnz = 0;
    int32_t d = 0;
    int value_bytes = 4;
    // CSR
    uint32_t *row_ptr = nullptr;
    int32_t *col = nullptr;
    void *val = nullptr;
    double *label = nullptr, *offset = nullptr, *weight = nullptr;
    // CSC + chunk metadata
    pml::Packed t;
    pml::RowSide rs;
    pml::SellSide sell;        // rows once more, sliced-ELL, when padding is cheap: the single-vector margin pass runs on it
    double *za = nullptr, *zb = nullptr;   // n each: sparse dots of row-pass variant 1 (zb lazily, Hessian-vector)
    double *carry_row = nullptr, *carry_row2 = nullptr;  // rs.n_chunks each
    // normalization
    double *factors = nullptr, *shifts = nullptr;
    int32_t intercept = -1;
    // per-evaluation scratch (allocated once at pack time)
    double *r = nullptr;           // n
    double *dcache = nullptr;      // n, lazily allocated on the first Hessian-vector call
    double *partials = nullptr;    // 2*MAX_PARTIALS
    double *carry_head = nullptr;  // n_chunks
    double *raw = nullptr;         // d+2
    double *vec = nullptr;         // 2*d: effective coefficient (pairs) buffer
    double *shift_partials = nullptr;  // 2*MAX_PARTIALS
    double *io_a = nullptr, *io_b = nullptr;  // d+2 each: staging of host coef / v / outputs
    double *io_n = nullptr;        // n, lazily: scores output staging
    double *row_loss = nullptr;    // n, lazily: per-row wt*l for per-entity objective values (pml_batch)
    bool want_row_loss = false;
    // batches with a mask (pml_batch_set_active): which chunks / row blocks still hold rows of an active entity
    unsigned char *row_chunk_active = nullptr, *col_chunk_active = nullptr, *row_block_active = nullptr;
    uint64_t cached_tag = 0;
    int cached_loss = -1;
    bool cache_valid = false;
    int lanes_per_row = 8;  // lanes-per-row variant used for paired (Hessian-vector, cold) passes, or the forced variant
    uint64_t version = 1;       // bumped by every change to rows / normalization / variant (invalidates vg_graph)
    pml_graph vg_graph;
    pml_graph hv_cold_graph, hv_cached_graph;  // Hessian-vector: margin pass included / curvature taken from dcache
    bool graph_failed = false;
    bool keep_column_order = false;  // batches: columns are compacted, not renumbered by frequency
    bool variant_forced = false;  // pml_dataset_set_margin_variant pinned one variant for every pass
    int row_grid = 1;       // CTAs of the last row-pass launch (= number of (F, P) partials)
    int64_t bytes = 0;
};
