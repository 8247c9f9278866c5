/*
 * avenir_tree.h -- C-ABI of libavenir_tree.so, the B200-native split-evaluation engine that sits
 * behind avenir's tree-building path (DecisionTreeBuilder / ClassPartitionGenerator / bagged
 * forests).
 *
 * The reference (kutieme/avenir) is pure JVM code and has NO native/FFI interface; the seams this
 * ABI replaces are Java object seams.  Each entry point below cites the reference code whose work
 * it takes over (paths relative to src/main/java/org/avenir/).  INTEGRATION.md shows the Panama
 * FFM / JNI binding a maintainer would add on the Java side.
 *
 * Conventions
 *   - plain C types, caller-owned host buffers, library-owned device memory behind an opaque handle;
 *   - every function returns 0 on success, a negative AVT_E_* code on failure; the message is
 *     available through avt_last_error();  no exception crosses the boundary;
 *   - one engine = one CUDA device + one stream; an engine is single-threaded, distinct engines are
 *     independent;  multi-GPU = one process (engine) per GPU joined by avt_comm_init_rank();
 *   - there is NO CPU fallback: without a CUDA device avt_create() fails.
 *
 * Data layout handed over by the host (encoded ONCE from the CSV rows, Appendix F of SURVEY.md):
 *   - columnar bin codes: for attribute a, column[a][r] = bin of row r (uint8 if n_bins[a] <= 256,
 *     else uint16).  Numeric: bin = #{grid points p_k < value}; categorical: index in `cardinality`;
 *   - class codes uint8[n_rows];
 *   - per tree: node id uint16[n_rows] (AVT_NODE_NONE = row not in any frontier node) and an
 *     optional bootstrap multiplicity uint8[n_rows].
 *
 * Developer switches (environment variables read once by avt_create; they select between equivalent code paths for
 * tests and A/B measurements and never change results; every rank of a multi-GPU job must run with the same values --
 * ranks that disagree on the exchange form fail with AVT_E_COMM after the bounded wait instead of hanging):
 *   AVT_SCAN_FAST=0   general scan kernel instead of the item-parallel one      AVT_LOOKBACK=0  ticket instead of look-back planning
 *   AVT_HIST_PACK=0   one byte per attribute instead of the bit-packed codes      AVT_HIST_JOINT=1 joint-code histogram kernel on packed schemas
 *   AVT_SUBTRACT=0|1  sibling subtraction off / on (default: on for >= 16 uint8 attributes)
 *   AVT_PDL=0         plain stream launches instead of programmatic dependent launch
 *   AVT_HIST_BUDGET_KB=n  rows/smem kernel: shared memory an attribute group may take (default 100)
 *   AVT_ROT_PF=n      forest kernel: node ids / multiplicities prefetched into L2 n trees ahead (default 4, 0 = off)
 *   AVT_HIST_SORTED=0|1|2  sorted-rows kernel off / only in place of the global-atomics fallback / whenever the forest kernel does not fit (default 2)
 *   AVT_PEER_PUSH=0   pull form of the NVLink peer exchange                        AVT_PEER_OFF=1  avt_peer_import is ignored (NCCL only)
 */
#ifndef AVENIR_TREE_H
#define AVENIR_TREE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AVT_OK 0
#define AVT_E_INVALID (-1)   /* bad argument / unsupported configuration */
#define AVT_E_CUDA (-2)      /* CUDA runtime error (message has the CUDA error string) */
#define AVT_E_STATE (-3)     /* call out of order (e.g. scan before histogram) */
#define AVT_E_NOMEM (-4)
#define AVT_E_COMM (-5)      /* NCCL error */
#define AVT_E_IO (-6)        /* tool level: file / parse error */
#define AVT_E_REFERENCE (-7) /* tool level: a configuration on which the reference itself fails */

#define AVT_NODE_NONE 0xFFFFu
#define AVT_MAX_CLASSES 16
#define AVT_MAX_SEG 8

#define AVT_CRIT_GINI 0
#define AVT_CRIT_ENTROPY 1
#define AVT_CRIT_HELLINGER 2
#define AVT_CRIT_CLASS_CONF 3

#define AVT_SPLIT_NUMERIC 0 /* segments are contiguous bin ranges given by upper bounds */
#define AVT_SPLIT_LUT 1     /* segments given by a per-bin uint8 segment table */
#define AVT_SPLIT_RANGES 2  /* segments are inclusive bin ranges [lo, hi] that MAY OVERLAP: the reference's multi-way numeric
                               splits (maxSplit > 2), whose last `le` predicate has no lower bound (tree/SplitManager.java:635-661)
                               and whose double twins drop every lower bound (:668-694, 840-843); SURVEY App. B-1 */

typedef struct avt_engine avt_engine;

/* One candidate split.  Splits of one attribute must be contiguous in the table; the table order
 * is the reference's enumeration order (tree/SplitManager.java:256-330,405-569;
 * explore/ClassPartitionGenerator.java:235-443), so index+1 within a node's candidate list is the
 * reference's running splitId (tree/DecisionTreeBuilder.java:299,325). */
typedef struct avt_split {
    int32_t attr;  /* attribute (column) index in the loaded dataset */
    int32_t kind;  /* AVT_SPLIT_NUMERIC | AVT_SPLIT_LUT */
    int32_t n_seg; /* number of segments, 2..AVT_MAX_SEG */
    int32_t off;   /* NUMERIC: index into `bounds` of n_seg-1 inclusive upper bin bounds;
                      LUT:     index into `seg_lut` of n_bins[attr] segment ids;
                      RANGES:  index into `bounds` of n_seg (lo, hi) pairs, hi < lo = empty segment.  Scan, eval and the
                               CPG sweep count a row once per segment it falls in; avt_grow_levels rejects tables with
                               RANGES splits (a row can have several children: the caller places them) */
} avt_split;

/* Result of the fused impurity scan for one frontier node. */
typedef struct avt_node_best {
    double min_av;      /* minimum weighted-average impurity over the node's candidate splits */
    double node_stat;   /* impurity of the node itself (InfoContentStat.processStat on its rows) */
    int64_t population; /* rows (weighted) in the node */
    int32_t best_split; /* lowest table index attaining min_av, -1 if the node has no candidate */
    int32_t n_ties;     /* number of candidate splits whose av compares == to min_av */
} avt_node_best;

/* Statistics of the children one (node, split) pair would create. */
typedef struct avt_split_eval {
    double av;                                        /* weighted average impurity of the split */
    double seg_stat[AVT_MAX_SEG];                     /* impurity per segment (0 if empty) */
    uint32_t seg_count[AVT_MAX_SEG][AVT_MAX_CLASSES]; /* class counts per segment */
} avt_split_eval;

/* Statistics of one ClassPartitionGenerator split (explore/ClassPartitionGenerator.java:514-553). */
typedef struct avt_cpg_stat {
    double stat;       /* SplitStat.processStat: weighted impurity / Hellinger / class-conf ratio */
    double split_info; /* SplitStat.getInfoContent: intrinsic information of the segment sizes */
    uint32_t seg_count[AVT_MAX_SEG][AVT_MAX_CLASSES];
} avt_cpg_stat;

typedef struct avt_stats {
    double hist_ms, scan_ms, eval_ms, repart_ms, comm_ms, h2d_ms; /* accumulated device time */
    int64_t hist_launches, scan_launches, eval_launches, repart_launches, comm_calls;
    int64_t hist_alg_bytes; /* algorithmic bytes of all histogram passes so far */
    int64_t repart_alg_bytes;
    int32_t last_hist_variant; /* 0 shared-memory privatised (lane = row), 1 global atomics (match_any), 2 forest kernel (hist_rot_kernel), 3 joint-code kernel (packed narrow schemas), 4 sorted-rows kernel (one tree, deep levels) */
    int32_t last_hist_replicas;
    int32_t last_hist_grid, last_hist_block, last_hist_smem;
    int32_t reserved0;
    double last_csv_h2d_ms, last_csv_index_ms, last_csv_encode_ms; /* last avt_dataset_load_csv: text upload, line index, per-row encode */
    double last_grow_ms; /* device time of the last avt_grow_levels call: CUDA events on the library's stream around everything
                          * the call enqueued, result copies included (host-side argument handling excluded) */
    double sample_ms;    /* accumulated device time of avt_forest_init_sampled (multiplicity generation + node-id reset) */
} avt_stats;

/* ---- engine ------------------------------------------------------------------------------- */

/* Creates an engine on CUDA device `device`.  Replaces nothing in the reference (which has no
 * device notion); it is the per-task state a BuilderMapper/BuilderReducer pair keeps in fields
 * (tree/DecisionTreeBuilder.java:102-131, 388-412). */
int avt_create(int device, avt_engine** out);
void avt_destroy(avt_engine* e);
/* Message of the last failure on `e` (or of the last failed avt_create when e == NULL). */
const char* avt_last_error(const avt_engine* e);

/* Uploads the once-encoded columnar data set.  Replaces the per-level re-read and re-parse of the
 * text rows in BuilderMapper.pathMapHelper (tree/DecisionTreeBuilder.java:266-320).
 * columns[a] is uint8[n_rows] when n_bins[a] <= 256, else uint16[n_rows].  The column and class buffers may be
 * host memory (pageable or page-locked) or memory of this engine's device (a caller that encodes on the GPU);
 * they are copied, the caller keeps ownership. */
int avt_dataset_load(avt_engine* e, int64_t n_rows, int32_t n_attrs, const int32_t* n_bins,
                     const void* const* columns, const uint8_t* class_codes, int32_t n_classes);

/* One feature attribute of a CSV row for avt_dataset_load_csv. */
#define AVT_CSV_INT 0         /* Integer.parseInt, bin = number of grid points < value */
#define AVT_CSV_DOUBLE 1      /* Double.parseDouble (surrounding blanks ignored), same binning */
#define AVT_CSV_CATEGORICAL 2 /* bin = index of the value in the attribute's value list */
typedef struct avt_csv_attr {
    int32_t ordinal; /* field index in the row (0-based, after `skip_fields` leading fields) */
    int32_t kind;    /* AVT_CSV_* */
    int32_t n;       /* grid points (numeric) or values (categorical); bins = n + 1 or n */
    int32_t off;     /* first entry in int_points / dbl_points / (value index in) cat_offsets */
} avt_csv_attr;

/* Why avt_dataset_load_csv rejected a row (bad_reason). */
#define AVT_CSV_BAD_SHORT_ROW 1 /* fewer fields than an attribute's or the class's ordinal needs */
#define AVT_CSV_BAD_INT 2       /* NumberFormatException of Integer.parseInt */
#define AVT_CSV_BAD_DOUBLE 3    /* NumberFormatException of Double.parseDouble */
#define AVT_CSV_BAD_CATEGORY 4  /* value not in the attribute's list ("no matching predicate") */
#define AVT_CSV_BAD_CLASS 5     /* class value not in class list */

/* Encodes CSV text into the columnar data set ON THE DEVICE and installs it like avt_dataset_load: the step
 * before the path.  Replaces, per row, the `split(fieldDelimRegex)` + Integer.parseInt /
 * Double.parseDouble + predicate-grid evaluation of BuilderMapper.pathMapHelper
 * (tree/DecisionTreeBuilder.java:268, 309-320) and the class extraction of BuilderReducer.reduce (:749).
 *   text/n_bytes   rows separated by '\n' (a '\r' before it is dropped; the last row may lack the '\n')
 *   delim          single-byte field delimiter (literal `field.delim.regex`)
 *   skip_fields    leading fields that are not part of the record (1 for rows tagged with their path)
 *   attrs          n_attrs attributes, ordinals strictly increasing; numeric grids in int_points / dbl_points
 *                  (ascending), categorical values as bytes cat_bytes[cat_offsets[off+k] .. cat_offsets[off+k+1])
 *   class values   n_class_values strings, class_bytes[class_offsets[k] .. class_offsets[k+1]); class code = k.
 *                  class_ordinal == -1: the rows carry no class field (tree/DataPartitioner.java's input); pass
 *                  n_class_values = 1, class_bytes / class_offsets may be NULL, every row gets class code 0
 *   class_first_row (nullable) [n_class_values]: first row each class value occurs in, -1 = never (lets the
 *                  caller rebuild the first-seen order the reference's HashMap insertion depends on)
 * On a row the reference would throw on, returns AVT_E_REFERENCE with *bad_row = the first such row and
 * *bad_reason = AVT_CSV_BAD_*; nothing is installed.  Values Double.parseDouble accepts but the device parser
 * cannot convert exactly (more than 15 significant digits, exponent beyond +-22, NaN/Infinity, hex) are
 * converted on the host with strtod and patched in; more than 65536 such values fail with AVT_E_INVALID. */
int avt_dataset_load_csv(avt_engine* e, const char* text, int64_t n_bytes, char delim, int32_t skip_fields,
                         int32_t n_attrs, const avt_csv_attr* attrs, const int32_t* int_points,
                         const double* dbl_points, const char* cat_bytes, const int32_t* cat_offsets,
                         int32_t class_ordinal, int32_t n_class_values, const char* class_bytes,
                         const int32_t* class_offsets, int64_t* n_rows, int64_t* class_first_row,
                         int64_t* bad_row, int32_t* bad_reason);

/* Copies the installed data set's codes back: column `attr` (uint8 or uint16 per row, as in avt_dataset_load;
 * attr == -1: the class codes).  For tests and for callers that keep a host copy. */
int avt_dataset_fetch_column(avt_engine* e, int32_t attr, void* out);

/* Physically partitions the ROWS OF THE TEXT the installed data set was encoded from (the last avt_dataset_load_csv on this
 * engine; the text and its line index stay on the device) by one attribute.  Replaces DataPartitioner's job
 * (tree/DataPartitioner.java:325-352): PartitionerMapper keys every row with AttributeSplitHandler.Split.getSegmentIndex of
 * its attribute value (util/AttributeSplitHandler.java:146-153, 190-204), the shuffle groups the rows by that key and
 * PartitionerReducer writes them out unchanged, one part file per reducer.
 *   attr             installed attribute (index into the attrs of avt_dataset_load_csv)
 *   segment_of_bin   [n_bins of attr] segment of every bin code, AVT_NODE_NONE = rows of this bin are dropped
 *   out_text         host buffer: the rows of segment 0 in input order, then segment 1, ...; every row is followed by
 *                    '\n' (TextOutputFormat), a '\r' before the source's line terminator is not part of the row
 *                    (LineRecordReader).  NULL: only the sizes are computed
 *   segment_bytes    [n_segments] bytes of each segment in out_text;  segment_rows (nullable) [n_segments] its rows
 * Fails with AVT_E_INVALID when out_capacity is smaller than the sum of segment_bytes (which are filled in). */
int avt_rows_partition(avt_engine* e, int32_t attr, const uint16_t* segment_of_bin, int32_t n_segments,
                       char* out_text, int64_t out_capacity, int64_t* segment_bytes, int64_t* segment_rows);

/* Starts `n_trees` trees over the loaded rows.  weights == NULL or weights[t] == NULL means
 * multiplicity 1 for every row (dtb.sub.sampling.strategy=none); otherwise weights[t][r] is the
 * number of times row r was drawn for tree t (explore/BaggingSampler.java:101-122 semantics).
 * node_ids == NULL or node_ids[t] == NULL puts every row in node 0 (the $root pass,
 * tree/DecisionTreeBuilder.java:248-258); otherwise node_ids[t][r] is the frontier node of row r
 * (AVT_NODE_NONE = dropped), as recovered from a path-tagged input file (:268-292). */
int avt_forest_init(avt_engine* e, int32_t n_trees, const uint8_t* const* weights,
                    const uint16_t* const* node_ids);

/* Starts `n_trees` trees whose first-pass sub-sampling is done ON THE DEVICE, for the rows loaded on this engine,
 * which are rows [row0, row0 + n_rows) of a file of `total_rows` rows (row0 = 0, total_rows = n_rows on one GPU; a
 * rank of a row-sharded job passes its shard's position, so that every shard count yields the same forest).
 * Tree t draws from a java.util.Random(seed + t) stream in file order (the reference's Math.random() is unseeded,
 * the seed is this library's addition).
 *   AVT_SAMPLE_WITH_REPLACE     param = batch size: per batch of `param` rows (the tail batch is shorter) m draws
 *                               (int)(nextDouble() * m); a row's multiplicity = draws that hit it.  This is
 *                               explore/BaggingSampler.java:101-122 (emitBatch) and the intended buffer bootstrap of
 *                               tree/DecisionTreeBuilder.java:214-229.  Fails if one exceeds 255.
 *   AVT_SAMPLE_WITHOUT_REPLACE  param = dtb.sub.sampling.rate: row kept (multiplicity 1) when
 *                               (int)(nextDouble() * 100) < param (tree/DecisionTreeBuilder.java:230-235).
 * Every row starts in node 0.  Same result as avt_forest_init with host-generated multiplicities. */
#define AVT_SAMPLE_WITH_REPLACE 1
#define AVT_SAMPLE_WITHOUT_REPLACE 2
int avt_forest_init_sampled(avt_engine* e, int32_t n_trees, int32_t mode, int64_t seed, int32_t param,
                            int64_t total_rows, int64_t row0);
/* Puts every row of every tree back into node 0 and keeps the trees' multiplicities: the level loop can be run again
 * on the same sample (asynchronous; ordered before the next call on this engine). */
int avt_forest_reset(avt_engine* e);
/* Copies back tree t's multiplicities (uint8[n_rows]; all 1 for a tree started without any). */
int avt_forest_weights(avt_engine* e, int32_t tree, uint8_t* out);

/* Installs the candidate-split table shared by all nodes. */
int avt_split_table_set(avt_engine* e, int32_t n_splits, const avt_split* splits,
                        const int32_t* bounds, int32_t n_bounds, const uint8_t* seg_lut,
                        int64_t n_lut);

/* Kernel (a): count histograms H[tree][node][attr][bin][class] for the current frontier, all
 * trees in one launch.  Replaces the mapper-side emission of one record per (row, split, true
 * predicate) plus the shuffle and the reducer-side counting (tree/DecisionTreeBuilder.java:
 * 323-352, 730-767; util/InfoContentStat.java:50-64).  nodes_per_tree[t] = frontier size of tree t.
 * With a communicator attached the histogram is all-reduced (sum) over the ranks.
 * hist_out (nullable) receives the [sum nodes][total_bins][n_classes] uint32 counts. */
int avt_level_histogram(avt_engine* e, const int32_t* nodes_per_tree, uint32_t* hist_out);

/* Kernel (b): fused gini/entropy scan + per-node arg-min.  Replaces BuilderReducer.expandTree's
 * per-parent split loop (tree/DecisionTreeBuilder.java:512-565) and InfoContentStat.processStat
 * (util/InfoContentStat.java:71-101).  cand_mask (nullable = every attribute) holds, per frontier
 * node, ceil(n_attrs/32) words with bit a set when attribute a is a candidate
 * (getSplitAttributes, :365-381).  class_order (nullable = 0..C-1) is the order in which class
 * terms are summed (HashMap order of the class strings).  out[node] is filled for
 * every frontier node. */
int avt_level_scan(avt_engine* e, int32_t criterion, const uint32_t* cand_mask,
                   const uint8_t* class_order, avt_node_best* out);

/* Per-split av values of one node from the last scan (NaN where the split's attribute is not a
 * candidate); the host needs them only to resolve ties in the reference's HashMap order. */
int avt_level_fetch_av(avt_engine* e, int32_t node, double* av_out);

/* Raw counts [total_bins][n_classes] of one frontier node of the last histogram; the host needs them
 * only to know which (split, segment) pairs received rows when it replays the reference's HashMap
 * insertion orders. */
int avt_level_fetch_hist(avt_engine* e, int32_t node, uint32_t* hist_out);

/* Children statistics for chosen (node, split) pairs (tree/DecisionTreeBuilder.java:575-611). */
int avt_level_eval(avt_engine* e, int32_t criterion, const uint8_t* class_order, int32_t n,
                   const int32_t* nodes, const int32_t* splits, avt_split_eval* out);

/* Kernel (c): row -> child node.  Replaces the re-keying of every row with its new path and the
 * next mapper's path filter (tree/DecisionTreeBuilder.java:333-349, 279-292).  For frontier node g
 * (global index over trees), node_attr[g] is the attribute of its selected split (-1: the node is
 * not expanded, its rows leave the frontier) and child_lut[g*lut_width + bin] the tree-local id of
 * the child that bin falls in (AVT_NODE_NONE: none). */
int avt_level_repartition(avt_engine* e, const int32_t* node_attr, const uint16_t* child_lut,
                          int32_t lut_width);

/* One record per expanded node of avt_grow_levels. */
typedef struct avt_grow_record {
    int32_t level, tree, node; /* node = tree-local id in that level's frontier */
    int32_t split;             /* chosen table index */
    int32_t n_ties;            /* >1: the reference would break the tie in HashMap order (App. B-7) */
    int32_t child[AVT_MAX_SEG]; /* tree-local id of each segment's child in the next frontier, -1 = empty */
    int64_t child_pop[AVT_MAX_SEG];
    double child_stat[AVT_MAX_SEG];
    double min_av;
} avt_grow_record;

/* The per-level driver loop on the device layout: `n_levels` times histogram -> scan -> children
 * statistics -> repartition for all trees at once, starting from the current node ids and frontier
 * sizes nodes_per_tree.  Replaces the hand-run loop of resource/detr.sh:34-54 (one Hadoop job + file
 * shuffle per level).  Every frontier node with rows and a candidate split is expanded (`stopped` is
 * advisory in the reference, App. B-6); ties go to the lowest table index and are reported in n_ties.
 * records (nullable) receives up to max_records records; *n_records the number produced;
 * nodes_per_tree is updated to the final frontier. */
int avt_grow_levels(avt_engine* e, int32_t n_levels, int32_t criterion, const uint8_t* class_order,
                    int32_t* nodes_per_tree, avt_grow_record* records, int64_t max_records,
                    int64_t* n_records);

/* avt_grow_levels with the reference's candidate-attribute strategy applied on the device (getSplitAttributes,
 * tree/DecisionTreeBuilder.java:365-381).  Every frontier node carries a candidate set of ceil(n_attrs/32) words:
 *   AVT_ATTRS_ALL           a child inherits its parent's set unchanged (SplitManager.getAllAttributes, tree/SplitManager.java:149-166;
 *                           with root_cand_mask = the dtb.custom.base.attributes subset);
 *   AVT_ATTRS_NOT_USED_YET  the reference's default, dtb.split.attribute.selection.strategy=notUsedYet: a child's set is its
 *                           parent's minus the attribute the parent was split on, i.e. the attributes not on its path
 *                           (SplitManager.getRemainingAttributes, tree/SplitManager.java:195-232).
 * root_cand_mask (nullable = every attribute) gives the sets of the nodes the call starts from, [sum nodes_per_tree][words].
 * A node whose set holds no attribute with a split is not expanded (the reference's mapper emits nothing for it). */
#define AVT_ATTRS_ALL 0
#define AVT_ATTRS_NOT_USED_YET 1
int avt_grow_levels_ex(avt_engine* e, int32_t n_levels, int32_t criterion, const uint8_t* class_order, int32_t strategy,
                       const uint32_t* root_cand_mask, int32_t* nodes_per_tree, avt_grow_record* records,
                       int64_t max_records, int64_t* n_records);

/* ClassPartitionGenerator sweep over every split of the installed table for the rows of frontier
 * node `node` (use 0 after avt_forest_init(1, NULL, NULL) for "all rows") of the last histogram.
 * Replaces PartitionGeneratorMapper.map + Combiner + Reducer (explore/ClassPartitionGenerator.java:
 * 200-230, 450-463, 514-592) and util/AttributeSplitStat.java:153-170,191-218,240-282,298-336.
 * class_order2 (Hellinger) = the two class codes in HashSet order. */
int avt_cpg_sweep(avt_engine* e, int32_t node, int32_t criterion, const uint8_t* class_order,
                  avt_cpg_stat* out);

/* Batch prediction with decision-path models on the loaded columns.  Replaces ModelPredictor.PredictorMapper.map
 * -> DecisionTreeModel.predictClassProb (tree/DecisionTreeModel.java:56-93: the first decision path, in list
 * order, whose predicates all hold) and, for n_models > 1, EnsemblePredictiveModel.predict
 * (model/EnsemblePredictiveModel.java:69-110: weighted vote, ties in the vote map's iteration order).
 * The host encodes each model as admission masks: masks of model m are total_bins x ceil(n_paths[m]/32) uint32
 * words, word w of (attribute a, bin b) at ((bin_off[a] + b) * words_m + w), bit p set when path 32*w+p admits
 * that bin; the models' blocks are concatenated in `masks`.  path_class holds, for every path of every model
 * (concatenated), the class code the path predicts (after cost-based thresholding, if any).  weights NULL = 1.0
 * each.  class_rank[c] = position of class c in the vote map's iteration order.  pred_out[r] = class code, or
 * 0xFF when some model has no matching path for row r (the reference dies with a NullPointerException). */
int avt_predict(avt_engine* e, int32_t n_models, const int32_t* n_paths, const uint32_t* masks,
                const uint8_t* path_class, const double* weights, int32_t n_classes,
                const uint8_t* class_rank, uint8_t* pred_out);

/* Copies back tree t's node ids (uint16[n_rows]). */
int avt_forest_node_ids(avt_engine* e, int32_t tree, uint16_t* out);
/* Copies back the class counts of every frontier node of the last histogram ([nodes][n_classes]). */
int avt_level_class_counts(avt_engine* e, uint32_t* out);

/* Device timer on the engine's stream (CUDA events): begin .. end -> elapsed milliseconds. */
int avt_timer_begin(avt_engine* e);
int avt_timer_end(avt_engine* e, float* ms);
int avt_synchronize(avt_engine* e);
int avt_get_stats(avt_engine* e, avt_stats* out);
int avt_reset_stats(avt_engine* e);
/* Turns the per-launch CUDA-event timing behind avt_stats on (default) or off (no event records between
 * the launches of a level loop). */
int avt_set_profiling(avt_engine* e, int32_t on);
/* Forces a histogram variant (-1 auto, 0 shared-memory by rows, 1 global atomics, 2 forest kernel over the rotated row-major
 * copy, 3 joint-code kernel for packed narrow schemas, 4 sorted-rows kernel for one tree) -- for tests/profiling. */
int avt_set_hist_variant(avt_engine* e, int32_t variant);

/* Allocation-free histogram.  NO COUNTERPART IN THE REFERENCE (and none in its Java object seams): it exists
 * for hosts that may not allocate device memory when they call -- a process whose GPU memory was divided
 * up at start-up, a thread on a latency budget that cannot take cudaMalloc/cudaFree's device-wide locks,
 * or code running while a stream capture is open elsewhere in the process.  Same counts as
 * avt_dataset_load + avt_forest_init(1 tree) + avt_level_histogram for one tree:
 *   hist_out[node][bin over all attributes][class], uint32, n_nodes x sum(n_bins) x n_classes.
 * columns / class_codes / node_ids are host buffers as in avt_dataset_load / avt_forest_init (node_ids NULL
 * = every row in node 0, AVT_NODE_NONE = row skipped).  All device-side storage is a fixed 8 MiB array
 * built into the library (one copy per GPU that loads it, kept until exit): no cudaMalloc, cudaFree,
 * stream or event creation happens.  Requests whose staged inputs plus counts exceed that array, or with
 * more than 64 attributes, fail with AVT_E_INVALID (avt_last_error(NULL) states needed vs available bytes).
 * The first call on each device looks up the array's device address: make it before any stream capture.
 * Calls on one GPU share the array; overlapping calls race -- serialise them.
 * The data is checked against the declared shape on the device: a node id >= n_nodes (other than AVT_NODE_NONE),
 * a bin code >= n_bins[a] or a class code >= n_classes fails the call with AVT_E_INVALID and leaves hist_out alone.
 * Side effects: the copies and the kernel run on the legacy default stream (which serialises with every blocking
 * stream of the process) and the call blocks until the counts are back; the caller's current device is restored. */
int avt_histogram_static(int device, int64_t n_rows, int32_t n_attrs, const int32_t* n_bins,
                         const void* const* columns, const uint8_t* class_codes, int32_t n_classes,
                         const uint16_t* node_ids, int32_t n_nodes, uint32_t* hist_out);

/* ---- multi-GPU: one process per GPU, rows sharded, histograms summed over NVLink ------------- */
/* Fills 128 bytes with an NCCL unique id (rank 0 calls it, the host broadcasts it). */
int avt_comm_unique_id(void* id128);
/* Joins the communicator.  From then on avt_level_histogram all-reduces H over the ranks: the one
 * exchange step of the path (the role of Hadoop's shuffle between BuilderMapper and
 * BuilderReducer, tree/DecisionTreeBuilder.java:80-93). */
int avt_comm_init_rank(avt_engine* e, int32_t n_ranks, int32_t rank, const void* id128);

/* NVLink peer exchange for avt_grow_levels (optional, ranks on one node): instead of an NCCL all-reduce per
 * level, every rank leaves its partial histogram in a buffer the others have mapped, and the scan kernel of
 * each rank sums the ranks' counts for its node straight out of peer memory -- the exchange step is fused
 * into kernel (b).  avt_peer_export allocates the exchange buffers (2 x 8 MiB + flags) and returns their
 * 64-byte CUDA IPC handle; the host gathers all ranks' handles (rank order) and passes the n_ranks x 64 bytes
 * to avt_peer_import.  Levels whose histogram exceeds the buffers fall back to the NCCL path (which then
 * needs avt_comm_init_rank).  Same role as avt_comm_*: Hadoop's shuffle between mapper and reducer.
 * One import per engine (a second one fails with AVT_E_STATE: the buffers still hold the first session's flags); an import
 * that fails part-way closes the mappings it had opened and leaves the engine without peers. */
int avt_peer_export(avt_engine* e, void* handle64);
int avt_peer_import(avt_engine* e, int32_t n_ranks, int32_t rank, const void* handles);

/* ---- tool level: the reference's own entry points -------------------------------------------- */
/* ToolRunner.run(new <tool>(), args): tool is "org.avenir.tree.DecisionTreeBuilder" or
 * "org.avenir.explore.ClassPartitionGenerator" (short names accepted); args are the Hadoop-style
 * arguments `-Dconf.path=<file.properties> [-Dkey=value ...] <inPath> <outPath>`
 * (tree/DecisionTreeBuilder.java:70-95,773-776; explore/ClassPartitionGenerator.java:64-91,611-614).
 * Returns 0 (success) or 1 (failure, message in errbuf) like the reference's exit status. */
int avt_tool_run(const char* tool, int argc, const char* const* argv, char* errbuf, size_t errlen);

#ifdef __cplusplus
}
#endif
#endif /* AVENIR_TREE_H */
