/* geobpe_b200 -- C ABI of the B200-native GeoBPE learn/induce hot path.
 *
 * The reference (shiningsunnyday/PT-BPE @ f032fa4) has NO FFI / plugin interface on this path:
 * it is pure Python.  The entry points below are therefore the boundary a
 * maintainer of the reference would bind with ctypes from foldingdiff/bpe.py; each one names the
 * reference Python code it replaces (file:line relative to the reference repo).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream); nothing here allocates,
 *     synchronises or owns memory -- buffers are owned by the caller (PyTorch tensors);
 *   - return value 0 = ok, anything else = failure, message in gb_last_error();
 *   - internal-coordinate input layout: `vals` is float64 [9][ld] type-major, row order
 *     0C:1N, N:CA, CA:C, phi, psi, omega, tau, CA:C:1N, C:1N:1CA (datasets.py:175-186),
 *     `offsets` int64 [S+1] is the CSR of backbones over residues;
 *   - "residue-major" slot order used by rbins / edges / class geometry:
 *       0 N:CA(r) 1 CA:C(r) 2 0C:1N(r) 3 tau(r) 4 CA:C:1N(r) 5 psi(r) 6 omega(r) 7 C:1N:1CA(r) 8 phi(r+1)
 *     (slots 0,1,3 of the first residue of a backbone are the NeRF init constants, tokenizer.py:74-77);
 *   - token arrays (one entry per live token, ordered by backbone then position):
 *       cls  u32  geometry class of the motif (medoid mode)        hsh u64  content hash (histogram mode)
 *       glue u32  packed glue bins to the right neighbour: omega | C:1N:1CA<<10 | phi<<20, bit31 = last token
 *       pos  u32  global residue index of the first residue         len u32  residues covered
 *       tid  u32  position of the token id in BPE._tokens
 */
#ifndef GEOBPE_B200_H
#define GEOBPE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GB_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
#define GB_LAST_FLAG 0x80000000u
#define GB_NOBIN 0x3FFu

const char* gb_last_error(void);
int gb_abi_version(void);
/* number of SMs of the current device (grid sizing; 148 on B200) */
int gb_sm_count(void);

/* ---- plumbing ------------------------------------------------------------------------------- */
/* bb[r] = backbone index of residue r (binary search in offsets).  Replaces the per-structure
 * Python loops `for ti, t in enumerate(self.tokenizers)` (bpe.py:229, 1440). */
int gb_residue_backbone(const int64_t* offsets, int64_t S, int64_t R, uint32_t* bb, void* stream);

/* ordered stream compaction: out_idx[0..total) = ascending i in [0,n) with (flags[i] & mask) != 0.
 * scratch: u32 [gb_select_scratch_elems(n)].  total is a device u32. */
int64_t gb_select_scratch_elems(int64_t n);
int gb_select_flagged(const uint8_t* flags, int64_t n, uint32_t mask, uint32_t* out_idx,
                      uint32_t* total, uint32_t* scratch, void* stream);
/* gb_select_flagged for callers that may not allocate (CUDA graph capture, serving loops): the block-sum scratch is a
 * file-scope device array inside the library (128 KB held by every GPU that loads it), so no scratch argument and no
 * cudaMalloc / cudaFree anywhere on the call.  No counterpart in the reference.
 * n above gb_select_flagged_static_max_n() (2^26 entries) is rejected with an error.  All calls share the
 * array: calls that overlap on the GPU (streams without a dependency between them) race; calls ordered on one stream
 * do not, and nothing in the array has to be cleared between calls. */
int64_t gb_select_flagged_static_max_n(void);
int gb_select_flagged_static(const uint8_t* flags, int64_t n, uint32_t mask, uint32_t* out_idx,
                             uint32_t* total, void* stream);
/* same, predicate keys[i] == key (filters the occurrence-set op log, bpe.py:1744 list(set)) */
int gb_select_eq_u64(const uint64_t* keys, int64_t n, uint64_t key, uint32_t* out_idx,
                     uint32_t* total, uint32_t* scratch, void* stream);

/* ---- K3: thresholds + value->bin (bpe.py:820-876 _init_thresholds, plotting.py:305-337
 *          save_histogram, bpe.py:1165-1189 get_ind, 1513-1526 quant_geo, 715-737 bond lengths) -- */
/* per column c: min / max / count over entries that are not NaN and not 0.0 (bpe.py:844), angle
 * columns (3..8) wrapped (v+2pi)%2pi first (plotting.py:318).  minmax f64 [9][2], cnt u64 [9]. */
int gb_col_minmax(const double* vals, int64_t R, int64_t ld, double* minmax, unsigned long long* cnt,
                  void* stream);
/* per column histogram against edges f64 [9][estride] with nb_host[c] bins (np.histogram semantics:
 * left-closed bins, last bin closed).  counts u64 [9][estride] is zeroed here. */
int gb_col_hist(const double* vals, int64_t R, int64_t ld, const double* edges, int estride,
                const int* nb_host, unsigned long long* counts, void* stream);
/* residue-major quantisation.  edges f64 [9][estride] in RESIDUE-MAJOR slot order, nb_host[9];
 * bond_mode 0: std_bonds (bond bins = 0), 1: free bonds strict, 2: free bonds clamped (strict=False,
 * bpe.py:726-729).  Outputs: rbins u16 [9][R] (GB_NOBIN = absent), glue u32 [R], sym u64 [R]
 * (1 + tau | CA:C:1N<<10 | psi<<20: the histogram-mode residue symbol), status i32 [S]
 * (glue angles and bonds: bit0 value below first bin, bit1 value in no bin -- get_ind's two ValueErrors;
 * motif-internal angles tau / CA:C:1N / psi: the same two in bits 2-3). */
int gb_quantize_residues(const double* vals, int64_t R, int64_t ld, const uint32_t* bb,
                         const int64_t* offsets, int64_t S, const double* edges, int estride,
                         const int* nb_host, int bond_mode, double init_nca, double init_cac,
                         double init_tau, uint16_t* rbins, uint32_t* glue, uint64_t* sym,
                         int32_t* status, void* stream);

/* ---- K1: adjacent-pair histogram + argmax (bpe.py:1431-1474 bin, 1192-1299 compute_geo_key,
 *          1796 peekitem, 2077-2138 recount) ------------------------------------------------------ */
/* table = open-addressing hash table of M (power of two) slots: keys u64, cnt u32, minpos u32 */
int gb_table_clear(uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin, int64_t M, void* stream);
/* medoid mode: key = cls[j]<<45 | cls[j+1]<<27 | glue27(j).  overflow (device i32) set if the table is full */
int gb_pair_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint64_t* tkeys,
                    uint32_t* tcnt, uint32_t* tmin, int64_t M, int32_t* overflow, void* stream);
/* histogram mode: key = polynomial hash over Z/2^64 (odd base P, natural wrap-around) of the concatenated window content
 * hsh[j] . gluesym(j) . hsh[j+1]; powtab u64 [>= 2*maxlen] holds P^m */
int gb_pair_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                    const uint64_t* powtab, uint64_t P, uint64_t* tkeys, uint32_t* tcnt,
                    uint32_t* tmin, int64_t M, int32_t* overflow, void* stream);
/* pair_count_* only fill keys and counts (one RED per distinct key per CTA).  The first position of every key tied
 * at the max is found afterwards by gb_first_positions_* (reads max / tie list from the device buffers written by
 * gb_table_argmax, so no host round trip lies between): tie_minpos[i] = lowest j whose pair (j,j+1) has key tie_keys[i] */
int gb_first_positions_m(const uint32_t* cls, const uint32_t* glue, int64_t T, const uint64_t* tkeys,
                         const uint32_t* tcnt, uint32_t* tmin, int64_t M, const uint32_t* out4,
                         const uint64_t* tie_keys, uint32_t* tie_minpos, int cap, void* stream);
int gb_first_positions_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                         const uint64_t* powtab, uint64_t P, const uint64_t* tkeys, const uint32_t* tcnt,
                         uint32_t* tmin, int64_t M, const uint32_t* out4, const uint64_t* tie_keys,
                         uint32_t* tie_minpos, int cap, void* stream);
/* generic: histogram of an explicit key array with first positions (residue-token interning, bpe.py:243-247) */
int gb_key_count(const uint64_t* keys, int64_t n, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                 int64_t M, int32_t* overflow, void* stream);
/* out u32 [4]: {max count, number of keys tied at max, distinct keys, 0}; ties (first `cap`) in
 * tie_keys / tie_minpos (unordered: the host applies the reference's json-string tie-break) */
int gb_table_argmax(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                    uint32_t* out4, uint64_t* tie_keys, uint32_t* tie_minpos, int cap, void* stream);
/* counts of Q given keys (recurring-key priority, bpe.py:2121-2127,2164-2166) */
int gb_table_lookup(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                    const uint64_t* qkeys, int Q, uint32_t* out_cnt, uint32_t* out_min, void* stream);
/* dump all (key,count,minpos) with count>0 compacted (unordered); total device u32 */
int gb_table_dump(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                  uint64_t* out_keys, uint32_t* out_cnt, uint32_t* out_min, uint32_t* total, void* stream);
/* multi-GPU: fold the table into a key-less direct-mapped counter array slots u32 [NS] (NS power of 2)
 * for the NCCL allreduce; and list the local keys that fall into slot `s` */
int gb_table_to_slots(const uint64_t* tkeys, const uint32_t* tcnt, int64_t M, uint32_t* slots,
                      int64_t NS, void* stream);

/* entries of this rank's table whose slot (same hash as gb_table_to_slots) is marked in slotmark u8 [NS] */
int gb_table_dump_marked(const uint64_t* tkeys, const uint32_t* tcnt, int64_t M, const uint8_t* slotmark,
                         int64_t NS, uint64_t* out_keys, uint32_t* out_cnt, uint32_t* total, int cap, void* stream);

/* ---- K1s: slot-sketch pair counting (same reference rows as K1: bpe.py:1431-1474 bin, 1796 peekitem,
 *      2077-2138 recount).  slots u32 [1 << log2ns], slot = (key * 0x9E3779B97F4A7C15) >> (64 - log2ns).
 *      variant bits: 1 = match.any warp aggregation, 2 = dynamic shared-memory hot stage, 4 = static hot list
 *      (hot_tags u32 [8192], 0xFFFFFFFF = unused), 8 = cache-streaming token loads, 16 = dry run (no REDs; timing
 *      probe only).  The caller zeroes `slots`. */
int gb_slot_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint32_t* slots, int log2ns,
                    const uint32_t* hot_tags, int variant, int ctas_per_sm, void* stream);
int gb_slot_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                    const uint64_t* powtab, uint64_t P, uint32_t* slots, int log2ns,
                    const uint32_t* hot_tags, int variant, int ctas_per_sm, void* stream);
/* sout u32 [4] = {max slot count, #candidates, #non-zero slots, 0}.  thr_host == 0: find the max and list the slots
 * holding it; thr_host > 0: list the slots with count >= thr_host (sout[0], sout[2], bmax of the thr_host == 0 call
 * are reused).  cand u32 [cap]; log2ns >= 12 */
int gb_slots_top(const uint32_t* slots, int log2ns, uint32_t thr_host, uint32_t* sout, uint32_t* cand, int cap,
                 uint32_t* bmax /* scratch u32 [(1 << log2ns) / 4096]: per-chunk maxima */, void* stream);
/* exact (key, count, first position) of every pair whose slot is in cand[0 .. *n_cand_dev) into the open-addressing
 * table (cleared by the caller).  overflow: 1 = table full, 2 = *n_cand_dev > cap (nothing counted); cap <= 4096 */
int gb_cand_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, int log2ns, const uint32_t* cand,
                    const uint32_t* n_cand_dev, int cap, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                    int64_t M, int32_t* overflow, int streaming, void* stream);
int gb_cand_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                    const uint64_t* powtab, uint64_t P, int log2ns, const uint32_t* cand,
                    const uint32_t* n_cand_dev, int cap, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                    int64_t M, int32_t* overflow, int streaming, void* stream);

/* ---- K2: merge (bpe.py:1895-2017) --------------------------------------------------------------- */
/* flags[j] bit0 = pair (j,j+1) has key `key` */
int gb_flag_pairs_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint64_t key, uint8_t* flags,
                    void* stream);
int gb_flag_pairs_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                    const uint64_t* powtab, uint64_t P, uint64_t key, uint8_t* flags, void* stream);
/* left-to-right overlap rule (bpe.py:1909-1914): within each maximal run of flagged pairs keep the
 * 1st, 3rd, ...: sets bit1 of flags on kept occurrences */
int gb_resolve_overlaps(uint8_t* flags, int64_t T, void* stream);
/* aux[occ[q]] = val[q] */
int gb_scatter_u32(const uint32_t* occ, const uint32_t* val, int64_t Q, uint32_t* aux, void* stream);
/* stream compaction of the token arrays.  A token survives unless its left neighbour pair was kept;
 * a kept left token j becomes the merged token: glue <- glue[j+1], len <- len[j]+len[j+1],
 * medoid mode: cls <- p2cls[aux[j]], tid <- p2tid[aux[j]]; histogram mode: hsh <- key, tid <- new_tid.
 * Any of cls/hsh may be NULL.  newT is a device u32.  scratch as gb_select_scratch_elems(T). */
int gb_merge_compact(const uint8_t* flags, int64_t T,
                     const uint32_t* cls, const uint64_t* hsh, const uint32_t* glue, const uint32_t* pos,
                     const uint32_t* len, const uint32_t* tid,
                     uint32_t* cls_o, uint64_t* hsh_o, uint32_t* glue_o, uint32_t* pos_o,
                     uint32_t* len_o, uint32_t* tid_o,
                     const uint32_t* aux, const uint32_t* p2cls, const uint32_t* p2tid,
                     uint64_t key, uint32_t new_tid, uint32_t* newT, uint32_t* scratch, void* stream);
/* merge-event log for the merge tree (data_structures.py:217-226): for kept occurrence q (token
 * kept[q]): ev_left[q]=pos[j], ev_right[q]=pos[j+1], ev_p[q]=aux? aux[j] : 0 */
int gb_write_events(const uint32_t* kept, int64_t Q, const uint32_t* count_dev /*nullable: device-side bound*/,
                    const uint32_t* pos, const uint32_t* aux,
                    uint32_t* ev_left, uint32_t* ev_right, uint32_t* ev_p, void* stream);
/* medoid mode: occurrence-set op log (SURVEY 7.2 hard part 1).  For kept occurrence q writes 4
 * records at log_key/log_pos[4q..4q+3]: remove left key, remove right key, add new left key, add
 * new right key (bpe.py:1924-1950,1992-2006), GB_EMPTY_KEY when the neighbour does not exist */
int gb_oplog_m(const uint32_t* kept, int64_t Q, const uint8_t* flags, const uint32_t* cls,
               const uint32_t* glue, const uint32_t* pos, int64_t T, const uint32_t* aux,
               const uint32_t* p2cls, uint64_t* log_key, uint32_t* log_pos, void* stream);
/* initial op-log segment = every adjacent pair in order (bpe.py:1441-1453): log_key[j], log_pos[j]=pos[j+1] */
int gb_pair_keys_m(const uint32_t* cls, const uint32_t* glue, const uint32_t* pos, int64_t T,
                   uint64_t* log_key, uint32_t* log_pos, void* stream);
/* histogram mode: exact check that every flagged occurrence has the same window bins as the window
 * at ref_r0 (guards the 61-bit content hash).  mismatches: device u32 counter */
int gb_verify_windows_h(const uint32_t* occ, int64_t Q, const uint32_t* pos, const uint32_t* len,
                        const uint16_t* rbins, int64_t R, uint32_t ref_r0, uint32_t nres,
                        uint32_t* mismatches, void* stream);

/* same check straight from the flags array; out2 u32 [2] = {mismatching windows, flagged pairs} */
int gb_verify_flagged_h(const uint8_t* flags, int64_t T, const uint32_t* pos, const uint32_t* len,
                        const uint16_t* rbins, int64_t R, uint32_t ref_r0, uint32_t nres, uint32_t* out2, void* stream);
/* packs {argmax out4, overflow, first nfirst ties: key, min position, pos, window residues} into res i64 [5+4*nfirst]
 * so the host reads one buffer per iteration; pos/len nullable (medoid mode needs no window) */
int gb_pack_result(const uint32_t* out4, const int32_t* overflow, const uint64_t* tie_keys,
                   const uint32_t* tie_minpos, const uint32_t* pos, const uint32_t* len, int nfirst,
                   const uint32_t* sout /* nullable: gb_slots_top's sout -> res[3] = sout[0] | sout[1] << 32 */,
                   int64_t* res, void* stream);

/* ---- device-driven histogram-mode learn loop (bpe.py:1792-2166 BPE.step with rmsd_key None, N iterations without a
 *      host round trip; csrc/k_hloop.cu).  args_host: int64 [gb_h_loop_nargs()] = device pointers and sizes in the order
 *      of the HA_* enum in k_hloop.cu (state block hs u32[32] + hs64 u64[2], ping-pong token arrays, flags + rmark u8,
 *      scan scratch, powtab, P, rbins, R, sketch buffers of gb_slot_count_h / gb_slots_top / gb_cand_count_h, tie list,
 *      JSON string scratch, merge-event log, per-step log, candidate-pair list, Nmax = upper bound of the array extent,
 *      streaming flag, compaction period).
 *      Per iteration the device picks the key with the largest pair count, breaks ties by the python string order of
 *      json.dumps(key, sort_keys=True) (bpe.py:1149,1469), merges its occurrences IN PLACE with the left-to-right overlap
 *      rule (the `len` arrays hold len | extra << 16 inside the loop: extra = dead entries up to the right neighbour) and
 *      appends (key, count, example window, event range) to the step log; the arrays are physically compacted every
 *      `period` iterations and, with compact_after, at the end (then len is plain again).  parity = which of the two
 *      token buffers is current on entry, *parity_out_host = on exit; step0 = iterations enqueued since the state block
 *      was last written by the host.  A condition the device cannot resolve sets hs[2] (halt code) and later
 *      iterations no-op. */
int gb_h_loop(const int64_t* args_host, int n_args, int nsteps, int parity, int step0, int compact_after,
              int* parity_out_host, void* stream);
int gb_h_loop_nargs(void);
/* sharded (multi-GPU) run: one iteration = parts 0..3 with a collective issued by the caller after parts 0-2
 * (allreduce-sum of the slot counters, allgather of the xbuf (key, count) lists, allreduce-max of the tie-string
 * buffer); see csrc/k_hloop_dist.cuh.  Compaction / state read-back go through gb_h_loop(nsteps = 0, compact_after = 1). */
int gb_h_loop_part(const int64_t* args_host, int n_args, int part, int parity, void* stream);
int64_t gb_h_loop_xbuf_elems(void);
int64_t gb_h_loop_sbuf_bytes(int stride);
/* peer-memory transport for the sharded run (replaces the three caller-issued collectives by loads over NVLink/NVSwitch
 * inside the loop's own kernels; csrc/k_hloop_dist.cuh).  Every rank allocates one symmetric block with gb_peer_alloc
 * ([slots u32 << log2ns | x u64 gb_h_loop_xbuf_elems() | sbuf gb_h_loop_sbuf_bytes() | flags u32 3 x gb_peer_max_world()],
 * each part 256-byte aligned), sends the 64-byte CUDA IPC handle to the other ranks (any transport) and maps theirs with
 * gb_peer_open.  args_host then carries a host table u64 [4][world] of all ranks' slots / x / sbuf / flags addresses
 * (HA_PEERS) plus local buffers for the reduced slot counters, the reduced tie strings and u32 [4] stage counters;
 * gb_h_loop_part then needs no collective between the parts.  A rank that sees no signal from a peer within 20 s halts
 * with code 9 instead of spinning forever. */
int gb_peer_alloc(int64_t bytes, void** ptr_out_host, unsigned char* handle64_out_host);
int gb_peer_open(const unsigned char* handle64_host, void** ptr_out_host);
int gb_peer_close(void* ptr);
int gb_peer_free(void* ptr);
int gb_peer_max_world(void);
/* whole compaction periods are replayed from captured CUDA graphs (default on) */
int gb_h_loop_use_graphs(int enable);
/* capture the graphs of both buffer parities now (warm-up) instead of at the first whole period */
int gb_h_loop_prepare(const int64_t* args_host, int n_args, void* stream);
int gb_h_loop_kernels_per_step(void);
/* slot counters inside the loop: 1 (default) = exact recount (k_slot_count) at the start of every compaction period and
 * upper bounds maintained from the merge events in between (hl_bump_list); 0 = recount every iteration */
int gb_h_loop_incremental(int enable);
/* 1 (default) = an inconclusive first candidate round is followed by a second one inside the loop (three more launches per
 * iteration that are no-ops ~999 times in 1000); 0 = such an iteration halts (code 2) and runs once on the host path
 * (faster steady state, but a halt drains the queue: measured worse end to end at 2 000 merges) */
int gb_h_loop_round2(int enable);
/* CUDA-event timing inside the loop: mode 1 = around every k_slot_count launch, 2 = around every first-round
 * hl_cand_count launch, 0 = off; run, synchronize, then read the per-launch ms */
int gb_h_loop_time_k1(int mode);
int gb_h_loop_k1_times(float* ms_host, int cap);

/* ---- device-driven learn loop, second design (csrc/k_hloop2.cu): the same iteration (bpe.py:1792-2166, rmsd_key None)
 *      inside ONE persistent cooperative kernel per call, with the pair counts maintained incrementally from the merge
 *      events the way the reference maintains _geo_dict / _priority_dict (bpe.py:1924-1950 decrements, 1992-2006 and
 *      2077-2138 increments / re-prioritisation) instead of being recounted: the slot counters of gb_slot_count_h stay
 *      exact (a destroyed pair decrements its slot, a created pair increments it); per "period" one streaming pass counts
 *      the keys of the candidate slots (count >= thr_num/16 of the top slot) exactly into an open-addressing table
 *      (key, count, example window) and lists their pairs; iterations then touch only the table, the pair list and the
 *      winner's occurrences.  A period ends on the device when the best exact count falls below the threshold, a slot
 *      outside the candidate set reaches it, the table / list fills or a quarter of the entries is dead (compaction).
 *      args_host: int64 [gb_h_loop2_nargs()] in the order of the HB_* enum (state block hs u32 [gb_h_loop2_state_words()]
 *      + hs64 u64 [16], ping-pong token arrays, flags + rmark u8, scan scratch, powtab, P, rbins, R, slot counters +
 *      chunk maxima + candidate list + cap + candidate bitmap u32 [NS/32], table tk/tc/tr0/tnres + M, pair list,
 *      occurrence list, JSON scratch + stride, merge-event log, per-step log, thr_num, streaming flag, CTAs per SM).
 *      Runs until hs[STEP] has advanced by nsteps or a halt code is set; final_compact != 0 leaves the arrays dense
 *      (len = plain residue counts) for host reads. */
int gb_h_loop2(const int64_t* args_host, int n_args, int nsteps, int final_compact, void* stream);
int gb_h_loop2_nargs(void);
int gb_h_loop2_state_words(void);
/* sharded run of the same loop (args: rbase = global index of the shard's first residue -- rbins then holds the rows of ALL
 * ranks --, world, rank, host table u64 [3][world] of every rank's delta / export / flag block inside its gb_peer_alloc'ed
 * symmetric block, dcap, xcap): slot counters, candidate bitmap and exact table are replicated; per iteration every rank
 * publishes the keys of the pairs its merges destroyed / created (bpe.py:1924-1950, 1992-2006) by STORING them into its
 * inbox slot of every rank's block over NVLink / NVSwitch from inside the kernel, and replays the other ranks' slots of its own block; per period the exact (key, count, example window) lists.  This
 * replaces the reference's single-process _geo_dict with one replicated priority structure -- "pair counts combined ...
 * merges applied locally" (north_star). */
int64_t gb_h_loop2_delta_bytes(int64_t dcap, int world);
int64_t gb_h_loop2_export_bytes(int64_t xcap, int world);
int64_t gb_h_loop2_flag_bytes(void);
int gb_h_loop2_max_world(void);
/* hsh[j] = content hash under base P of token j's symbol sequence (residue symbols sym, glue symbols glue_res + 1) --
 * the value the merges build with (H(L) P + g) P^|R| + H(R); used to re-key a run after a detected 64-bit collision */
int gb_rehash_tokens(const uint64_t* sym, const uint32_t* glue_res, const uint32_t* pos, const uint32_t* len, int64_t T,
                     uint64_t P, uint64_t* hsh, void* stream);

/* ---- token-stream export: Tokenizer.tokenize (tokenizer.py:379-392) + BPE.quantize (bpe.py:928-956) of every
 *      backbone in one pass.  out i64 [4 T - 3 S]: token j of backbone b writes at 4 j - 3 b its motif id tid[j] and, unless it
 *      is the backbone's last token, V + off_omega + bin(omega), V + off_phi + bin(phi), V + off_cn + bin(C:1N:1CA) -- the
 *      reference's order DIHEDRAL, DIHEDRAL, BOND_ANGLE behind a residue-aligned token; off_* = cum_bin_count(type). */
int gb_token_stream(const uint32_t* tid, const uint32_t* glue, const uint32_t* pos, const uint32_t* bb, int64_t T,
                    uint32_t V, uint32_t off_omega, uint32_t off_phi, uint32_t off_cn, int64_t* out, void* stream);

/* ---- reconstruction metrics for every backbone (bin/encode.py:341-359 computes them per reference tokenizer through
 *      Tokenizer.compute_coords + ProteinChain.rmsd): current value of every residue-major slot (bin centres centres f64
 *      [9][cstride]; medoid mode: the live token's class geometry cg f64 [ncls][maxres][9], NaN = absent, glue at token ends from
 *      the bins; cg == NULL in histogram mode), pristine values in the same layout, full-chain NeRF of both (coords f64
 *      [3 R][3]) and the Kabsch backbone RMSD per backbone (rmsd f64 [S]).  tok_of u32 [R], slots_* f64 [9][R] are scratch. */
int gb_reconstruction(const uint16_t* rbins, int64_t R, const double* centres, int cstride, const uint32_t* pos,
                      const uint32_t* len, const uint32_t* cls, int64_t T, const double* cg, int maxres,
                      const double* vals, int64_t ld, const uint32_t* bb, const int64_t* offsets, int64_t S,
                      double i_nca, double i_cac, double i_tau, uint32_t* tok_of, double* slots_cur, double* slots_orig,
                      double* coords_cur, double* coords_orig, double* rmsd, void* stream);
/* lDDT-C-alpha of every backbone between two coordinate sets laid out like gb_reconstruction's (f64 [R][3 atoms][3], atom 1 =
 * CA): what bin/encode.py:351-358 and plotting.py:436 take from esm's ProteinChain.lddt_ca(...).mean() per reference tokenizer.
 * esm is absent from the reference tree and this image; its published compute_lddt is restated (float32, eps 1e-10, inclusion
 * radius `cutoff` = 15 A in the reference's call, thresholds 0.5 / 1 / 2 / 4 A, per-residue scores averaged).  lddt f64 [S]. */
int gb_lddt_ca(const double* coords_pred, const double* coords_true, const int64_t* offsets, int64_t S, double cutoff,
               double* lddt, void* stream);

/* ---- K7: induce merge replay on the device (bpe.py:1116-1139 BPE.tokenize's loop over the learned merges + step_helper
 *      1316-1425, for a whole batch of held-out backbones at once).  Token arrays are the medoid-mode SoA in the lazily
 *      compacted layout (ls = len | extra << 16, 0 = dead entry; dense on entry).  Merge m (host arrays, launch order =
 *      vocabulary order, merges that cannot occur left out): packed pair key, K medoids with A = bonds + 1 atoms each
 *      (coordinates f64 at med + mmed_off[m]), class / token id of medoid p at p2cls / p2tid [mp2_off[m] + p].  Per merge:
 *      occurrences flagged in one pass, each assigned to its nearest medoid (NeRF from the pristine coordinates `vals`,
 *      Kabsch RMSD, first minimum), overlap rule, in-place merge, (left, right, medoid) appended to the event log.
 *      st u32 [8] zeroed by the caller except st[2] = first free event slot; rep_start u32 [n_merges + 1] with
 *      rep_start[0] = st[2]; rep_cnt u32 [n_merges] receives the events of every merge.  Nothing is read back.
 *      cw u32 [N]: scratch for the flag pass's stream words (class | 14-bit hash of the glue bins), filled here. */
int gb_induce_replay(uint32_t* cls, uint32_t* glue, const uint32_t* pos, uint32_t* ls, uint32_t* tid, uint8_t* flags,
                     uint8_t* rmark, uint32_t* aux, uint32_t* cw, int64_t N, const double* vals, int64_t ld, const uint32_t* bb,
                     const int64_t* offsets, const double* f3tab, int n_merges, const uint64_t* mkey_host,
                     const int* mK_host, const int* mA_host, const int* mnres_host, const int64_t* mmed_off_host,
                     const int64_t* mp2_off_host, const double* med, const uint32_t* p2cls, const uint32_t* p2tid,
                     uint32_t* occ_j, uint32_t* occ_r, uint32_t* ev_left, uint32_t* ev_right, uint32_t* ev_p,
                     int64_t evcap, uint32_t* st, uint32_t* rep_start, uint32_t* rep_cnt, void* stream);
int gb_induce_max_atoms(void);

/* ---- K4/K5/K6: NeRF windows, Kabsch RMSD, medoid assignment (tokenizer.py:347-363, nerf.py:84-210,
 *      angles_and_coords.py:238-317, algo.py:8-65,170-181, bpe.py:646-657) -------------------------- */
/* coords f64 [Q][A][3] for windows starting at residue r0[q] spanning nres residues; A = atoms kept
 * (3*nres+1, or 3*nres when the window ends a backbone).  If the window starts a backbone the first
 * N:CA, CA:C, tau are f3tab[f3cls[q]] when f3cls != NULL and f3cls[q] != 0xFFFFFFFF, else (d_nca,d_cac,d_tau)
 * (the reference keeps them in Tokenizer._init_*; tokenizer.py:131-167).
 * bondc (nullable) f64 [3][nbmax]: bin centres for 0C:1N / N:CA / CA:C with brow u16 [3][R] bins --
 * when given, bond lengths are taken quantised (induce, bpe.py:1061,1085). */
int gb_window_coords(const double* vals, int64_t R, int64_t ld, const uint32_t* bb, const int64_t* offsets,
                     const uint32_t* r0, int64_t Q, int nres, int A,
                     const uint32_t* f3cls, const double* f3tab, double d_nca, double d_cac, double d_tau,
                     const uint16_t* rbins, const double* bond_centres, int bc_stride,
                     double* coords, void* stream);
/* assign[q] = argmin_m rmsd(P_q, M_m) (first index on ties, np.argmin), best[q] = that rmsd */
int gb_kabsch_assign(const double* P, int64_t Q, int A, const double* Mq, int K, uint32_t* assign,
                     double* best, void* stream);
/* D f32 [N][N] all-pairs Kabsch RMSD (float64 math, rounded to float32 like algo.py:170) */
int gb_rmsd_matrix(const double* P, int N, int A, float* D, void* stream);
/* the k-medoids iterations of algo.py:184-211 on D (one CTA): med_init u32 [k] = the host's rng.choice draw, med_out u32 [k],
 * st u32 [4] = {status, iterations, float bits of the last total_shift, -}.  status 1 = a cluster came out empty, which
 * draws rng.integers in the reference (algo.py:199): repeat the call's iterations on the host from med_init (the kernel
 * uses no random numbers).  Every numpy operation is restated bit for bit: first-minimum argmin, float32 pairwise
 * row sums (gb_pairwise_sum_host = the same routine compiled for the host, pinned against numpy on the CPU), float32
 * total_shift.  N <= gb_pam_max_n(), k <= gb_pam_max_k(), else an error (run the host loop). */
int gb_kmedoids_pam(const float* D, int N, int k, const uint32_t* med_init, int max_iter, uint32_t* med_out,
                    uint32_t* st, void* stream);
int gb_pam_max_n(void);
int gb_pam_max_k(void);
float gb_pairwise_sum_host(const float* v, int n);

#ifdef __cplusplus
}
#endif
#endif
