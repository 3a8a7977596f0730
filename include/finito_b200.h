/*
 * finito_b200 — C ABI of the B200-native finimizer hot path.
 *
 * The reference (ElenaBiagi/Finito @ ae6e702) has no FFI; the path sits behind its C++ object API
 * and its two CLI commands.  Each entry point below replaces one reference interface, batched so
 * that one call keeps a whole GPU busy.  Reference citations are file:line under the reference root.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every buffer it passes in, the library owns
 *     device memory behind the opaque handles;
 *   - every function returns fin_status; on failure fin_last_error() (thread local) says why;
 *   - no exception crosses this boundary, no CPU fallback exists: without a usable CUDA device every
 *     compute entry point returns FIN_ERR_CUDA;
 *   - a handle is bound to one CUDA device; one host thread drives one handle at a time.
 */
#ifndef FINITO_B200_H
#define FINITO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum fin_status {
    FIN_OK = 0,
    FIN_ERR_ARG = 1,     /* bad argument (null pointer, capacity too small, unknown type string) */
    FIN_ERR_IO = 2,      /* file missing / truncated / not writable */
    FIN_ERR_FORMAT = 3,  /* file is not what the reference would have written (variant, version) */
    FIN_ERR_CUDA = 4,    /* CUDA runtime failure or no device */
    FIN_ERR_DOMAIN = 5,  /* input outside the parity domain: non-ACGT base, unitig < k ... */
    FIN_ERR_LIMIT = 6    /* index too large for the index width this build supports */
} fin_status;

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* fin_last_error(void);

/* ABI version of this header (bumped on any signature change). */
int fin_abi_version(void);

/* Number of visible CUDA devices (0 when there is none; never fails). */
int fin_device_count(void);

/* ------------------------------------------------------------------------------------------------
 * FinimizerIndex  (include/FinimizerIndex.hh:108-115 members, :209-241 load)
 * ---------------------------------------------------------------------------------------------- */
typedef struct fin_index fin_index;

typedef struct fin_index_info {
    int64_t k;             /* sbwt->get_k() */
    int64_t n_nodes;       /* sbwt->number_of_subsets() */
    int64_t n_kmers;       /* sbwt->number_of_kmers() */
    int64_t n_unitigs;     /* unitigs.ends.size() */
    int64_t text_len;      /* unitigs.concat.size() */
    int64_t n_finimizers;  /* global_offsets.size() */
    int64_t device_bytes;  /* bytes of HBM held by the device layout */
    int64_t device;        /* CUDA device ordinal */
} fin_index_info;

/* FinimizerIndex::load(prefix) (FinimizerIndex.hh:209-241): reads <prefix>.{LCS,FBV,O,packed_unitigs,
 * unitig_endpoints,Ustart}.sdsl and <prefix>.sbwt, re-lays them out for the device and uploads them
 * to CUDA device `device`. */
fin_status fin_index_load(const char* prefix, int device, fin_index** out);
void fin_index_free(fin_index* index);
fin_status fin_index_info_get(const fin_index* index, fin_index_info* out);

/* ------------------------------------------------------------------------------------------------
 * search-fmin  (include/search_fmin.hh:43-72 = per read: FinimizerIndex::search(read),
 * FinimizerIndex::search(reverse complement), merge; FinimizerIndex.hh:119-185; common.hh:78-186)
 * ---------------------------------------------------------------------------------------------- */
typedef struct fin_search_counts {
    int64_t found_fwd;     /* sum of QueryResult::n_found over the forward searches   (search_fmin.hh:66) */
    int64_t found_rc;      /* ... over the reverse-complement searches                (search_fmin.hh:67) */
    int64_t positive;      /* merged k-mer positions with a hit                       (search_fmin.hh:58) */
    int64_t n_queries;     /* k-mer positions answered = sum max(0, len-k+1)          (search_fmin.hh:69) */
} fin_search_counts;

/* Number of k-mer positions of a batch: sum over reads of max(0, len - k + 1).  Pure host arithmetic. */
uint64_t fin_count_kmers(const uint64_t* read_offsets, uint64_t n_reads, int64_t k);

/* One batch of reads, HOST buffers.  `seqs` holds the reads back to back (ASCII, upper or lower case
 * ACGT, no separators), read i = seqs[read_offsets[i] .. read_offsets[i+1]).  `out_pairs` receives,
 * read after read, one (unitig id, offset) pair per k-mer position, (-1,-1) where neither strand has
 * the k-mer — exactly the numbers search-fmin prints (search_fmin.hh:54-65).  `out_capacity` is in
 * pairs and must be >= fin_count_kmers().  The call copies the reads to the device, runs the
 * kernels and copies the hits back (pipelined in sub-batches); it returns when `out_pairs` is complete.
 * A non-ACGT character gives FIN_ERR_DOMAIN (the reference has undefined behaviour there,
 * common.hh:108-111).  Pinned host buffers (fin_host_alloc) make the copies asynchronous. */
fin_status fin_search_batch(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                            int64_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts);

/* FinimizerIndex::search(query) for every read of the batch (include/FinimizerIndex.hh:119-185): the forward
 * strand only, no reverse-complement search and no merge — the call the reference's own tests make
 * (src/tests.cpp:86-101).  counts->found_fwd is the sum of QueryResult::n_found. */
fin_status fin_search_forward(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                              int64_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts);

/* Same, 2 x int32 per k-mer position (halves the device->host traffic).  FIN_ERR_LIMIT if a unitig id
 * or offset of this index cannot be represented. */
fin_status fin_search_batch32(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                              int32_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts);

/* Same computation on buffers that already live on the index's device (used to time the kernels
 * without the PCIe copies).  d_seqs: ASCII reads; d_read_offsets: n_reads+1 offsets; d_out: 2 x int32
 * (hit_bytes == 4) or 2 x int64 (hit_bytes == 8) per k-mer position; total_kmers = fin_count_kmers of
 * the batch (the caller knows it from the host copy of the offsets).  Synchronous: returns after the
 * kernels have finished.  kernel_ms (may be NULL) receives the CUDA-event time of the kernels. */
fin_status fin_search_device(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                             uint64_t n_bases, uint64_t total_kmers, void* d_out, int hit_bytes,
                             fin_search_counts* counts, float* kernel_ms);

/* fin_search_device for callers that cannot tolerate an allocation (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Same arguments, same answers and same errors as fin_search_device, but the call
 * issues no cudaMalloc / cudaFree / allocating library routine: every scratch array of the search (2-bit reads, output
 * offsets, found mask, strand flags, accumulators) lives in ONE static __device__ array of 8 MiB inside the library,
 * which every GPU that loads the library holds until the process exits.  Limits per call: n_reads <=
 * FIN_NOALLOC_MAX_READS and n_bases <= FIN_NOALLOC_MAX_BASES, otherwise FIN_ERR_LIMIT and nothing is launched.
 * All calls (all handles, all threads) on one GPU share the array: it keeps the last call's contents (the call clears
 * what must start at zero) and calls that overlap on the same GPU race - serialise them.  Runs on the default stream
 * and returns after the kernels have finished.  (The first call on a device resolves the array's device address with
 * cudaGetSymbolAddress; a call with n_reads == 0 can serve as that set-up.) */
#define FIN_NOALLOC_MAX_READS 65536ull
#define FIN_NOALLOC_MAX_BASES 16777216ull
fin_status fin_search_device_noalloc(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                                     uint64_t n_bases, uint64_t total_kmers, void* d_out, int hit_bytes,
                                     fin_search_counts* counts);

/* ---- run-length hits -----------------------------------------------------------------------------
 * 98 % of the found k-mers of a read come out of walk_in_unitigs (include/FinimizerIndex.hh:47-102): consecutive
 * k-mer positions at consecutive offsets of one unitig.  A run says that for one maximal stretch the search produced in
 * one go: merged k-mer positions [first_kmer, first_kmer + len) of its read map to `unitig`; position first_kmer + d
 * has offset first_offset + d, or first_offset - d when FIN_RUN_RC is set (a reverse-complement hit: the unitig runs
 * against the read, include/search_fmin.hh:54-60).  The runs of a read are sorted by first_kmer and never overlap;
 * every position outside a run is (-1,-1).  fin_expand_runs turns them back into the table fin_search_batch* fills,
 * bit for bit; at 150 bp / 1 % error a read has ~3 runs (48 bytes) instead of 120 pairs (960 bytes). */
typedef struct fin_run {
    uint32_t first_kmer;   /* merged k-mer position of the run's first entry within its read */
    uint32_t len_rc;       /* length | FIN_RUN_RC */
    uint32_t unitig;
    uint32_t first_offset; /* offset of position first_kmer in the unitig */
} fin_run;
#define FIN_RUN_RC 0x80000000u

/* fin_search_batch32 with the answer as runs.  run_offsets (n_reads + 1 entries): the runs of read i are
 * runs[run_offsets[i] .. run_offsets[i+1]).  *n_runs receives the number of runs of the batch; if it exceeds
 * runs_capacity the call returns FIN_ERR_ARG, `runs` is incomplete and *n_runs says how much room a retry needs
 * (fin_count_kmers() always suffices: a run holds at least one k-mer position). */
fin_status fin_search_batch_runs(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                                 uint64_t* run_offsets, fin_run* runs, uint64_t runs_capacity, uint64_t* n_runs,
                                 fin_search_counts* counts);

/* Same on DEVICE buffers (d_seqs, d_read_offsets as for fin_search_device; d_run_offsets: n_reads + 1 u64;
 * d_runs: runs_capacity records).  Synchronous. */
fin_status fin_search_device_runs(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                                  uint64_t n_bases, uint64_t* d_run_offsets, fin_run* d_runs, uint64_t runs_capacity,
                                  uint64_t* n_runs, fin_search_counts* counts, float* kernel_ms);

/* Runs -> the (unitig, offset) table of fin_search_batch (hit_bytes 8) / fin_search_batch32 (hit_bytes 4) on the
 * HOST, with n_threads worker threads (<= 1: the calling thread only).  Pure host arithmetic over the caller's
 * arrays: no device, no handle. */
fin_status fin_expand_runs(const uint64_t* read_offsets, uint64_t n_reads, int64_t k, const uint64_t* run_offsets,
                           const fin_run* runs, void* out_pairs, uint64_t out_capacity, int hit_bytes, int n_threads);

/* Access counters of the last search call on this handle, the ones SURVEY.md section 8(d) builds the
 * algorithmic-bytes figure from.  They are only tallied while counting is enabled (fin_search_set_counting,
 * or FIN_COUNTERS=1 in the environment at load time): the counting kernel variant is slower. */
typedef struct fin_access_counters {
    int64_t n_rank, n_lcs_reads, n_ustart_reads, n_anchor, n_anchor_branch, n_walk_words, n_bases, n_kmers;
    int64_t n_filter_lookups; /* s-mer table lookups of the strand prefilter (not part of the reference's algorithm) */
    /* n_rank and n_lcs_reads tally what the REFERENCE's loops access.  These two say how many of them this library never
     * issues: the shrink loop and the contraction loops are evaluated in closed form (one drop instead of one per character). */
    int64_t n_lcs_skipped, n_rank_skipped;
} fin_access_counters;
fin_status fin_search_access_counters(const fin_index* index, fin_access_counters* out);
fin_status fin_search_set_counting(fin_index* index, int enabled);

/* CUDA-event times of the last fin_search_device call: the whole pipeline, the preparation kernels
 * (2-bit packing, k-mer counts, output offsets) and the streaming-search kernel alone. */
typedef struct fin_search_timing {
    float pipeline_ms, prepare_ms, search_kernel_ms;
} fin_search_timing;
fin_status fin_search_last_timing(const fin_index* index, fin_search_timing* out);

/* Number of kernel launches issued by this library on this thread's handles since load (bench.py's
 * gpu_launches claim is read from here). */
uint64_t fin_launch_count(void);

/* Measurement aid: GB/s of data-independent random reads of `bytes_per_access` (32, 64 or 128: aligned runs of 32-byte sectors) over a table of
 * `table_bytes` (use >= 4 GiB so that the 126 MB L2 does not matter), best of `repeats`.  The HBM random-access roofline
 * bench.py reports beside the copy peak. */
fin_status fin_probe_random_sectors(int device, uint64_t table_bytes, int repeats, int bytes_per_access, double* gb_per_s);

/* Pinned host memory for the batch buffers (cudaHostAlloc / cudaFreeHost). */
fin_status fin_host_alloc(size_t bytes, void** out);
void fin_host_free(void* p);

/* ------------------------------------------------------------------------------------------------
 * build-fmin  (include/build_fmin.hh:302-401 driver; FinimizerIndex.hh:273-389 builder;
 * lcs_basic_parallel_algorithm.hpp:52-120; PackedStrings.hh:104-137)
 * ---------------------------------------------------------------------------------------------- */

/* build-fmin -i sbwt_path -u unitig_path -o out_prefix --type type -t t [--lcs lcs_path].
 *   sbwt_path : plain-matrix .sbwt as written by `sbwt build` (with the variant string).
 *   unitig_path: FASTA/FASTQ (optionally .gz) or a .txt list of such files (build_fmin.hh:338).
 *   type      : "rarest" | "shortest" | "verify" (build_fmin.hh:367-371).  rarest with t != 1 fails
 *               with the reference's message (build_fmin.hh:245-247).
 *   lcs_path  : NULL or "" = compute the LCS on the device and write <out_prefix>.LCS.sdsl
 *               (build_fmin.hh:373-380).
 * rarest writes the seven index files byte-identical to the reference's; every type appends
 * "<t>,<stats>\n" to <out_prefix>_stats.txt (build_fmin.hh:386-395).  stats_out (may be NULL)
 * receives <stats>. */
fin_status fin_build(const char* sbwt_path, const char* unitig_path, const char* out_prefix, const char* type,
                     int64_t t, const char* lcs_path, int device, char* stats_out, size_t stats_cap);

/* Wall-clock stages of the last fin_build call on this thread (milliseconds): the figures behind bench.py's
 * "build-fmin unitig bp/sec".  stream_kernel_ms is the CUDA-event time of build_stream_kernel. */
typedef struct fin_build_timing {
    float load_sbwt_ms, lcs_ms, read_unitigs_ms, prepare_ms, stream_merge_ms, stream_kernel_ms, write_ms, total_ms;
    int64_t unitig_bases;
} fin_build_timing;
fin_status fin_build_last_timing(fin_build_timing* out);

/* LCS array only (lcs_basic_parallel_algorithm.hpp:52-120) as one byte per SBWT column. */
fin_status fin_lcs(const char* sbwt_path, int with_variant_header, int device, uint8_t* out, uint64_t out_capacity,
                   uint64_t* n_nodes);

/* ------------------------------------------------------------------------------------------------
 * sbwt build (the producer of build-fmin's input): the plain-matrix SBWT of the k-mers of a
 * FASTA/FASTQ(.gz) file (or .txt list of files), constructed on the device and written in the format
 * `sbwt build --variant plain-matrix` writes (SBWT/src/CLI/sbwt_build.cpp:30-160, SBWT/include/sbwt/SBWT.hh:395-424):
 * same node set and order as the reference constructors (SBWT/include/sbwt/NodeBOSSInMemoryConstructor.hh:98-188,
 * kmc_construct.hh:42-238), no reverse complements added, k-mer prefix precalc table of precalc_k characters
 * (the tool's default is 8; clamped to k).  2 <= k <= 64 in this build (FIN_ERR_LIMIT above). */
fin_status fin_sbwt_build(const char* unitig_path, int64_t k, const char* out_sbwt_path, int64_t precalc_k, int device);

/* Wall-clock stages of the last fin_sbwt_build call on this thread (milliseconds). */
typedef struct fin_sbwt_build_timing {
    float read_ms, keys_ms, sort_ms, nodes_ms, write_ms, total_ms;
    int64_t bases, n_nodes, n_kmers;
} fin_sbwt_build_timing;
fin_status fin_sbwt_build_last_timing(fin_sbwt_build_timing* out);

/* Disjoint spectrum-preserving string set of the k-mers of a FASTA/FASTQ(.gz) file, built on the device: the stand-in
 * for the unitig tools the reference's README runs in front of `sbwt build` (ggcat, unitig_flipper; README.md:22-40).
 * Every canonical k-mer (a k-mer and its reverse complement are one) is kept at its first occurrence; the output FASTA
 * holds the maximal stretches of consecutive kept k-mers, each >= k long, reverse-complemented at random when flip != 0.
 * Valid input for fin_sbwt_build + fin_build (every k-mer exactly once); not maximal unitigs. */
typedef struct fin_spss_info {
    int64_t n_strings, bases, n_kmers, input_bases;
} fin_spss_info;
fin_status fin_spss_build(const char* in_path, int64_t k, const char* out_fasta_path, int flip, uint64_t seed, int device,
                          fin_spss_info* info);

/* ------------------------------------------------------------------------------------------------
 * build-fmin --type rarest sharded over GPUs: one process (or thread) per GPU, every rank opens the same
 * inputs.  Replaces the single loop over the permuted unitigs of FinimizerIndexBuilder's constructor
 * (include/FinimizerIndex.hh:292-299) by `world` partial streams whose (finimizer column, overwrite-rule key)
 * tuples meet at the owner of the column.  Call order on every rank:
 *   fin_builder_open   -> loads, LCS (computed on the device when lcs_path is NULL/""; rank 0 writes the file),
 *                         unitig permutation + Ustart (replicated), chunk plan, this rank's share
 *   fin_builder_stream -> add_sequence (FinimizerIndex.hh:321-389) over this rank's chunks; *n_tuples = number of
 *                         columns this rank produced a key for
 *   fin_builder_export -> d_tuples (DEVICE, n_tuples x {u64 column, u64 key}, ascending by column);
 *                         counts_per_rank (HOST, world entries): how many of them rank g owns, in order
 *   ... the caller exchanges the tuples: ONE all-to-all (torch.distributed / NCCL in finito_b200/distributed.py) ...
 *   fin_builder_merge  -> folds the received tuples (DEVICE) into this rank's slice of columns
 *                         [rank * slice_columns, (rank+1) * slice_columns); *n_slice_fmin = finimizers in the slice
 *   fin_builder_slice  -> d_bits (DEVICE, slice_words u64: fmin bits of the slice), d_offsets (DEVICE,
 *                         n_slice_fmin u32: global_offsets of the slice in column order, FinimizerIndex.hh:301-306)
 *   ... the caller gathers the slices on rank 0 ...
 *   fin_builder_finish -> rank 0 only: all slices in rank order (HOST) -> the seven index files + the stats line
 * With world == 1 export/exchange are skipped: open, stream, merge(NULL, 0), slice, finish  (what fin_build does).
 * ---------------------------------------------------------------------------------------------- */
typedef struct fin_builder fin_builder;

typedef struct fin_builder_info {
    int64_t n_nodes, slice_columns, slice_words, text_len, n_unitigs;
    int64_t n_chunks, chunk_begin, chunk_end; /* the chunk plan and this rank's share [begin, end) */
    int64_t bases_streamed;                   /* bases this rank's share streams (incl. the 2k warm-up per chunk) */
    int32_t rank, world;
    float stream_ms;                          /* CUDA-event time of build_stream_kernel (after fin_builder_stream) */
} fin_builder_info;

fin_status fin_builder_open(const char* sbwt_path, const char* unitig_path, const char* lcs_path, const char* out_prefix,
                            int device, int rank, int world, fin_builder** out);
void fin_builder_free(fin_builder* b);
fin_status fin_builder_info_get(const fin_builder* b, fin_builder_info* out);
fin_status fin_builder_stream(fin_builder* b, uint64_t* n_tuples);
fin_status fin_builder_export(fin_builder* b, void* d_tuples, uint64_t capacity_tuples, uint64_t* counts_per_rank);
fin_status fin_builder_merge(fin_builder* b, const void* d_tuples, uint64_t n_tuples, uint64_t* n_slice_fmin);
fin_status fin_builder_slice(const fin_builder* b, void* d_bits, void* d_offsets);
fin_status fin_builder_finish(fin_builder* b, const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all,
                              uint64_t n_offsets, const char* out_prefix);

/* ------------------------------------------------------------------------------------------------
 * build-fmin --type shortest | verify sharded over GPUs (include/build_fmin.hh:95-214): one process (or thread) per GPU,
 * every rank opens the same inputs.  The reference streams the unitigs in file order with one shared `fmin_found` bit
 * vector; its insertion rule has an order-free closed form (a window inserts iff no window with the same finimizer column
 * that ends at or after position k-1 of its unitig comes earlier in the file), so rank r streams a share of the unitigs and
 * the ranks meet twice.  Call order on every rank:
 *   fin_stats_open        -> loads, LCS (rank 0 writes <out_prefix>.LCS.sdsl when out_prefix is not NULL/""), this rank's share
 *   fin_stats_first_pass  -> shortest: table[column] = first such window of this rank's share; *n_table = its entries
 *                            (n_nodes; 0 for verify, which needs no table)
 *   fin_stats_table(out)  -> DEVICE u32[n_table]  ... the caller reduces the tables with MIN over the ranks (one all-reduce) ...
 *   fin_stats_table(in)   -> the reduced table
 *   fin_stats_second_pass -> this rank's keys (len, freq, column), *n_keys of them
 *   fin_stats_keys        -> DEVICE u64[n_keys]  ... the caller concatenates the keys of all ranks (one all-gather) ...
 *   fin_stats_finish      -> any rank: sort + unique + sums over ALL keys (DEVICE) = the reference's std::set; stats_out
 *                            receives "<n>,<sum freq>,<avg freq>,<avg len>,<n k-mers>" (common.hh:188-206); write_line != 0
 *                            appends "<t>,<stats>\n" to <out_prefix>_stats.txt (build_fmin.hh:386-395)
 * world == 1: open, first_pass, second_pass, keys, finish  (what fin_build does for these types).
 * ---------------------------------------------------------------------------------------------- */
typedef struct fin_stats fin_stats;
fin_status fin_stats_open(const char* sbwt_path, const char* unitig_path, const char* lcs_path, const char* out_prefix, const char* type,
                          int64_t t, int device, int rank, int world, fin_stats** out);
void fin_stats_free(fin_stats* s);
fin_status fin_stats_first_pass(fin_stats* s, uint64_t* n_table);
fin_status fin_stats_table(fin_stats* s, void* d_table, int set);
fin_status fin_stats_second_pass(fin_stats* s, uint64_t* n_keys);
fin_status fin_stats_keys(const fin_stats* s, void* d_keys);
fin_status fin_stats_finish(fin_stats* s, const void* d_keys_all, uint64_t n_keys_all, int write_line, char* stats_out, size_t stats_cap);

#ifdef __cplusplus
}
#endif
#endif /* FINITO_B200_H */
