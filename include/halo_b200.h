/* halo_b200 — C ABI of the B200-native MSM / IPA / NTT path for rasmus-kirk/halo.
 *
 * These are the symbols a `halo-b200-sys` Rust crate would bind (INTEGRATION.md shows the extern block and the
 * call-site patches).  Conventions, all taken from the reference:
 *   - field element  = 4 x u64 little-endian limbs, Montgomery form, R = 2^256, value < modulus
 *                      (crates/group/src/wrappers.rs:603-616)
 *   - affine point   = WrappedPoint {x:[u64;4], y:[u64;4]} (#[repr(C)], wrappers.rs:592-597); identity = (0,0)
 *                      (wrappers.rs:91-93)
 *   - projective pt  = Jacobian (X,Y,Z), 12 x u64, as ark_ec Projective (crates/plonk/src/frontend/ivc/mod.rs:40-45);
 *                      identity has Z = 0.  Any representative of the group element may be returned.
 *   - curve: 0 = Pallas, 1 = Vesta.   field: 0 = halo Fp (Pallas scalars), 1 = halo Fq (Vesta scalars)
 *                      (crates/group/src/lib.rs:8-9).  A curve's scalars live in field == curve.
 * Every function returns 0 on success or a negative HALO_E* code; nothing aborts.  Pointers are HOST pointers
 * unless the name ends in _dev.  The library keeps no caller pointer after returning.  One process drives one GPU;
 * N GPUs are N such processes joined by the halo_dist_* calls below.
 */
#ifndef HALO_B200_H
#define HALO_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HALO_OK 0
#define HALO_EINVAL (-1)   /* bad argument (the reference would assert!/panic!) */
#define HALO_ECUDA (-2)    /* CUDA runtime error; see halo_last_error() */
#define HALO_ENOINIT (-3)  /* halo_init not called */
#define HALO_EPARAMS (-4)  /* public parameters missing or too small (pp.rs:27-32, 63-82) */
#define HALO_EREJECT (-5)  /* a verifier rejected (anyhow::ensure! sites: pcdl.rs:494,547-550,580; acc.rs:154,207-210) */

/* ---- runtime ------------------------------------------------------------------------------------------ */
/* Binds the library to CUDA device `device` of this process (-1 = current device) and creates its streams.
 * Idempotent. */
int halo_init(int device);
int halo_shutdown(void);
const char* halo_last_error(void);
/* Use the caller's CUDA stream (cudaStream_t) for all subsequent work, so that the caller can bracket calls with
 * its own events; NULL restores the library's stream. */
int halo_set_stream(void* cuda_stream);
int halo_device_synchronize(void);

/* ---- public parameters  ≙ PublicParams::{new,set_pp,get_pp}  (crates/group/src/pp.rs:26-94) ------------- */
/* Uploads n generators + S + H once per curve; idempotent and monotone like set_pp: a later call with a larger
 * n than the resident set fails with HALO_EPARAMS, a smaller or equal one is a no-op.  Every point is checked to
 * be on the curve (wrappers.rs:606); failure -> HALO_EINVAL. */
int halo_pp_load(int curve, const uint64_t* gs_xy, size_t n, const uint64_t s_xy[8], const uint64_t h_xy[8]);
/* number of resident generators (0 = none) */
/* The reference's own parameter files (crates/group/.precompute/<curve>/gs-NN.bin and sh.bin, bincode 2 `standard()`
 * varint encoding of Vec<WrappedPoint> / (WrappedPoint, WrappedPoint); crates/group/src/pp.rs:36-53).
 * halo_pp_decode_block decodes one file on the host (with_length = 1 for gs-NN.bin, 0 for sh.bin with max_points = 2)
 * into x,y limb octets, taking at most max_points as `.take(m)` does; no device needed.
 * halo_pp_load_blobs ≙ PublicParams::new(n) (pp.rs:26-61): the first n generators of the blocks in order plus S, H,
 * made resident like halo_pp_load (same on-curve check, same idempotent / monotone behaviour).
 * HALO_EINVAL on a malformed block (the reference panics), HALO_EPARAMS when the blocks hold fewer than n points. */
int halo_pp_decode_block(const uint8_t* bytes, size_t len, int with_length, size_t max_points, uint64_t* out_xy,
                         size_t* n_points);
int halo_pp_load_blobs(int curve, const uint8_t* const* blocks, const size_t* block_lens, size_t n_blocks,
                       const uint8_t* sh, size_t sh_len, size_t n);
size_t halo_pp_len(int curve);
/* drop the resident set (tests only; the reference's OnceLock cannot be reset) */
int halo_pp_reset(int curve);

/* ---- MSM  ≙ point_dot_affine / pedersen::commit / pcdl::commit  (group.rs:48-50, pedersen.rs:7-27, pcdl.rs:275) */
/* out = sum_{i<k} scalars[i] * B[i];  B = bases_xy (k points) or, when bases_xy == NULL, resident Gs[base_off..]. */
int halo_msm(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k,
             uint64_t out_xyz[12]);
/* b independent MSMs over the same resident bases Gs[base_off .. base_off+k): scalars is b x k x 4 words.
 * ≙ the 16 commits of a prover round (crates/plonk/src/plonk/protocol.rs:114,263). */
int halo_msm_batch(int curve, size_t base_off, const uint64_t* scalars, size_t k, size_t b, uint64_t* out_xyz);
/* projective bases  ≙ point_dot (group.rs:53-56): normalize_batch + MSM */
int halo_msm_proj(int curve, const uint64_t* bases_xyz, const uint64_t* scalars, size_t k, uint64_t out_xyz[12]);
/* sum of n Jacobian points (host arithmetic): combines the per-GPU partial sums of a point-range-sharded MSM */
int halo_point_sum(int curve, const uint64_t* pts_xyz, size_t n, uint64_t out_xyz[12]);
/* same as halo_msm with inputs already in device memory (d_bases may be NULL for resident Gs) */
int halo_msm_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k,
                 uint64_t out_xyz[12]);
/* batched variant over device memory: b scalar vectors of k elements against the same k bases */
int halo_msm_batch_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k, size_t b,
                       uint64_t* out_xyz);
/* experiment knobs (0 = default): key 0 = fixed-base table for the first IPA generator fold (1 = never, 2 = for every opening size; default: openings
 * over the whole parameter set of >= 2^16 generators), key 1 = 1 disables the one-launch path of MSMs with k <= 2^13,
 * key 2 = 1 forces the literal 255-bit ladder in the IPA fold instead of the endomorphism split,
 * key 3 = log2 of the vector length below which the IPA stops folding generators (-1 never, 0 default),
 * key 4 = 1 disables the radix-4 (two-at-once) generator folds of the IPA,
 * key 5 = 1 makes the PLONK prover always take the 16-coset coefficient-division path for the quotient,
 * key 6 = 1 makes the PLONK prover transform and commit all-zero columns like any other (no sparse-column path),
 * key 7 = 1 prints a per-round phase trace of halo_pcdl_open to stderr,
 * key 9 = 1 disables the point-range pipelining of halo_msm over host buffers (upload overlapped with accumulation),
 * key 10 = number of point ranges of that pipeline (2..16; default 4),
 * key 8 = 64 / 128 / 256: bucket accumulation of the general MSM pipeline by batched affine additions with that many
 *        positions per thread (opt-in: measured slower than the default XYZZ task path on these 255-bit fields),
 * key 13 = 1 runs the two openings of a PLONK proof one after the other instead of side by side,
 * key 14 = 1 disables the fixed-base table of the resident generators (commit path, first L / R pair of an opening),
 * key 15 = 1 computes the witness commitments of a PLONK proof after round 0 instead of beside it */
int halo_set_tuning(int key, int value);
/* window width override for experiments (0 = automatic) */
int halo_msm_set_window(int c);

/* ---- PC_DL and AS_DL  (crates/accumulation/src/pcdl.rs, acc.rs) ---------------------------------------------
 * Flat layouts in u64 words, L = lg n = log2(d + 1):
 *   proof    = Ls[12 L] Rs[12 L] U[12] c[4] hiding[1] C_bar[12] w_prime[4]              (24 L + 33 words, EvalProof)
 *   instance = C[12] d[1] z[4] v[4] proof                                               (24 L + 54 words, Instance)
 * Points are Jacobian, scalars Montgomery; hiding = 0 leaves C_bar / w_prime zero (Option::None).
 * The Fiat-Shamir transcript (crates/poseidon) runs inside the library on the host. */
size_t halo_instance_words(unsigned lg_n);
size_t halo_proof_words(unsigned lg_n);
/* ≙ pcdl::commit(p, d, w) (pcdl.rs:275-287) against the resident generators; asserts -> HALO_EINVAL */
int halo_pcdl_commit(int curve, const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w_or_null, uint64_t out_xyz[12]);
/* ≙ pcdl::chunked_commit(p, d, w, chunk_size) (pcdl.rs:294-314): one commit per chunk of p's coefficients (trailing zeros
 * stripped as DensePolynomial does), each against Gs[0..chunk) with the same optional w.  out_xyz must hold
 * ceil(len / chunk_size) points; *out_count (may be NULL) receives the number written.  chunk_size == 0 or a chunk
 * longer than n -> HALO_EINVAL (slice::chunks / pedersen.rs:14-19 panic). */
int halo_pcdl_chunked_commit(int curve, const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w_or_null, size_t chunk_size,
                              uint64_t* out_xyz, size_t* out_count);
/* Explicit set-up of the fixed-base table the first generator fold of an opening of size n uses (2 x 136 x n/2 affine
 * points: 9.1 GB at n = 2^20, ~0.25 s).  halo_pcdl_open builds it lazily for openings over the whole parameter set of
 * >= 2^16 generators; a one-shot caller can pay for it here, outside its timed path.  Freed by halo_pp_reset /
 * halo_shutdown.  HALO_ECUDA when the device has no room (openings then use the ladder fold). */
int halo_ipa_fold_table_build(int curve, size_t n);
/* ≙ pcdl::open / open_without_eval (pcdl.rs:326-473).  v_or_null == NULL: v = p(z) is computed (open) and written to
 * out_v_or_null.  Hiding (w_or_null != NULL) consumes the caller's randomness: hide_q = the d coefficients of the
 * uniform polynomial q (pcdl.rs:348), hide_wbar = w_bar (pcdl.rs:352). */
int halo_pcdl_open(int curve, const uint64_t* coeffs, size_t len, const uint64_t C_xyz[12], size_t d, const uint64_t z[4],
                   const uint64_t* v_or_null, const uint64_t* w_or_null, const uint64_t* hide_q, const uint64_t* hide_wbar,
                   uint64_t* out_proof, uint64_t* out_v_or_null);
/* ≙ pcdl::succinct_check (pcdl.rs:483-554): HALO_OK and (h as its L+1 challenges, U), or HALO_EREJECT */
int halo_pcdl_succinct_check(int curve, const uint64_t* instance, uint64_t* out_xis_or_null, uint64_t* out_U_or_null);
/* ≙ pcdl::check (pcdl.rs:563-583) */
int halo_pcdl_check(int curve, const uint64_t* instance);
/* ≙ acc::prover / verifier / decider (acc.rs:178-217) over m instances of equal d; the accumulator is an instance */
int halo_acc_prover(int curve, const uint64_t* instances, size_t m, uint64_t* out_acc);
int halo_acc_verifier(int curve, const uint64_t* instances, size_t m, const uint64_t* acc);
int halo_acc_decider(int curve, const uint64_t* acc);
/* commit(h) for h(X) = prod (1 + xi_{L-i} X^{2^i}) expanded on the device  ≙ pcdl.rs:577 */
int halo_hpoly_commit(int curve, const uint64_t* xis, size_t lg_n, uint64_t out_xyz[12]);
/* coefficients of sum_i alphas[i] * h_i(X)  ≙ AccumulatedHPolys::get_poly (acc.rs:86-92); xis: m x (L+1) x 4 */
int halo_acc_hpoly(int field, const uint64_t* xis, const uint64_t* alphas, size_t m, size_t lg_n, uint64_t* out_coeffs);
/* ≙ HPoly::eval (pcdl.rs:221-233), host */
int halo_hpoly_eval(int field, const uint64_t* xis, size_t lg_n, const uint64_t z[4], uint64_t out[4]);

/* ---- wire formats  ≙ the derived CanonicalSerialize / CanonicalDeserialize (ark-serialize 0.5, Compress::Yes) of Instance,
 * EvalProof and Accumulator (crates/accumulation/src/pcdl.rs:26,159; acc.rs:22) as qs.bin is written
 * (crates/accumulation/src/main.rs:59-91).  Host code; no device and no halo_init needed.
 *   scalar: 32 bytes little-endian canonical integer.   point: affine form, 32 bytes LE canonical x + 1 flag byte
 *   (bit 7: y is the larger root, bit 6: infinity).   usize: u64 LE.   Vec: u64 LE length + items.   Option: 0 / 1 + item.
 * Instance = C d z v Ls Rs U c C_bar? w_prime?; an Accumulator is its instance q, byte for byte.
 * Deserialisation validates what ark's does (scalars below the modulus, x below the modulus with x^3 + 5 a square, not
 * both flags) and additionally requires what the flat layout assumes: Ls and Rs of lg(d + 1) points, C_bar and w_prime
 * both present or both absent.  Malformed input -> HALO_EINVAL. */
int halo_point_compress(int curve, const uint64_t xyz[12], uint8_t out[33]);
int halo_point_decompress(int curve, const uint8_t in[33], uint64_t out_xyz[12]);
int halo_scalar_serialize(int field, const uint64_t s[4], uint8_t out[32]);
int halo_scalar_deserialize(int field, const uint8_t in[32], uint64_t out[4]);
/* *written (may be NULL) always receives the size needed: 253 + 66 lg n bytes with hiding, 188 + 66 lg n without */
int halo_instance_serialize(int curve, const uint64_t* instance, uint8_t* out, size_t cap, size_t* written);
/* out_instance: halo_instance_words(lg n) words for the d found in the bytes; *consumed (may be NULL) = bytes read */
int halo_instance_deserialize(int curve, const uint8_t* bytes, size_t len, uint64_t* out_instance, size_t cap_words,
                              size_t* consumed);

/* ---- multi-GPU: one process per GPU, the library owns an NCCL communicator ------------------
 * The reference is a single process with rayon; a host that wants N GPUs runs N ranks (threads of a launcher, MPI
 * ranks, torchrun workers), each of which calls halo_init(device) and then halo_dist_init.  Rank 0 obtains the 128-byte
 * NCCL id with halo_dist_unique_id and ships it to the others over any channel it has (a file, a socket, MPI_Bcast).
 * NCCL is bound with dlopen("libnccl.so.2") on first use, so single-GPU callers never need it.
 * Every halo_dist_* call is COLLECTIVE: all ranks call it, in the same order, and all receive the same result.
 *   MSM        ≙ point_dot_affine / commit sharded by POINT RANGE: a rank runs the whole Pippenger pipeline over its
 *              range down to its window sums, which are all-gathered from device memory and combined on every rank.
 *              halo_dist_msm takes the rank's own slice (k_local may differ between ranks and may be 0);
 *              halo_dist_msm_global takes the WHOLE problem on every rank (the SPMD drop-in for group.rs:48-50) and
 *              slices it internally; bases_xy == NULL uses the resident generators Gs[base_off ..].
 *   AS_DL      ≙ acc::prover / verifier / decider sharded by INSTANCE: succinct checks of instances rank, rank + N, ...
 *              on each rank's host cores, h(X) = sum alpha^i h_i(X) as per-rank partial coefficient vectors all-gathered
 *              and added on the device, the decider's commit(h) as a point-range-sharded MSM.  The transcript and the
 *              opening are sequential and run replicated.  Every rank passes all m instances.
 *   NTT        batches sharded by POLYNOMIAL; a single transform stays on one GPU.  Every rank passes the whole batch
 *              and receives the whole result. */
int halo_dist_unique_id(uint8_t out[128]);
int halo_dist_init(int rank, int world, const uint8_t nccl_unique_id[128]);
int halo_dist_shutdown(void);
/* HALO_ENOINIT when no communicator exists; rank / world may be NULL */
int halo_dist_info(int* rank, int* world);
int halo_dist_msm(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k_local,
                  uint64_t out_xyz[12]);
int halo_dist_msm_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k_local,
                      uint64_t out_xyz[12]);
int halo_dist_msm_global(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k,
                         uint64_t out_xyz[12]);
/* b <= 64 MSMs over the same local base range, device-resident scalars (b x k_local x 4 words) */
int halo_dist_msm_batch_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k_local, size_t b,
                            uint64_t* out_xyz);
int halo_dist_acc_prover(int curve, const uint64_t* instances, size_t m, uint64_t* out_acc);
int halo_dist_acc_verifier(int curve, const uint64_t* instances, size_t m, const uint64_t* acc);
int halo_dist_acc_decider(int curve, const uint64_t* acc);
int halo_dist_ntt(int field, uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null);
/* Poseidon on the device: `count` INDEPENDENT sponges, one GPU thread each.  Sponge k absorbs
 * inputs[k * len .. (k + 1) * len) (Montgomery elements of `field`, 4 words each) and squeezes one element into out[k] --
 * InnerSponge::new / absorb / squeeze of crates/poseidon/src/inner_sponge.rs:11-115 (Kimchi parameters: x^7, 55 full
 * rounds, width 3, rate 2; constants of crates/group/src/poseidon_consts.rs).  A single Fiat-Shamir transcript is a
 * serial chain and stays on the host inside the library; this entry is for batches (hash-to-field over a column, Merkle
 * layers, many transcripts at once).  len = 0 squeezes the empty sponge.  The _dev variant takes device pointers and
 * only enqueues on the library stream. */
int halo_poseidon_hash_batch(int field, const uint64_t* inputs, size_t len, size_t count, uint64_t* out);
int halo_poseidon_hash_batch_dev(int field, const void* d_inputs, size_t len, size_t count, void* d_out);
/* pcdl::open (crates/accumulation/src/pcdl.rs:326-473) over the GPUs of the communicator.  Collective; every rank passes
 * the SAME arguments (those of halo_pcdl_open) and receives the same proof, equal to halo_pcdl_open's.  The rounds above
 * the fold-free tail run on interleaved slices (rank r owns the indices j * world + r of the generator, coefficient and
 * power vectors, so every fold pairs two elements of the same rank): L / R are point-range-sharded MSMs, the two dot
 * products are summed over the ranks; at <= 2^15 generators the slices are all-gathered and the latency-bound tail runs
 * replicated.  Needs a power-of-two world size and n >= 2^17, otherwise it is halo_pcdl_open on every rank.
 * halo_dist_acc_prover's opening is sharded the same way. */
int halo_dist_pcdl_open(int curve, const uint64_t* coeffs, size_t len, const uint64_t C_xyz[12], size_t d, const uint64_t z[4],
                        const uint64_t* v_or_null, const uint64_t* w_or_null, const uint64_t* hide_q, const uint64_t* hide_wbar,
                        uint64_t* out_proof, uint64_t* out_v_or_null);
/* ONE PLONK proof over the GPUs of the communicator (SURVEY section 8(e) / BASELINE config 4 at 2, 4, 8 GPUs;
 * PlonkProof::naive_prover, crates/plonk/src/plonk/protocol.rs:64-355 -- the reference is one process with rayon).
 * Collective.  polys / w_evals / public_inputs / acc_prev are READ ON RANK 0 ONLY (the other ranks may pass NULL): the
 * columns cross the host link once and are broadcast over NVLink.  Every rank runs the same transcript; the commitments
 * (protocol.rs:114,263) are point-range-sharded MSMs, the cosets of the quotient (protocol.rs:170-268) are divided
 * among the ranks and all-gathered (when the world size divides 8), the two openings (protocol.rs:295-296) run on
 * ranks 0 and 1, the serial pieces (grand product, evaluations, accumulation) are replicated.  Every rank receives the
 * proof, identical to halo_plonk_prove's.  lg_n, n_public (and the handle's circuit) must agree on all ranks; the
 * _circuit variant needs halo_plonk_circuit_load of the same circuit on every rank. */
int halo_dist_plonk_prove(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                          size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof);
int halo_dist_plonk_prove_circuit(int handle, const uint64_t* w_polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                                  size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof);

/* ---- NTT / iNTT  ≙ Evals::from_poly* / Evals::interpolate*  (crates/group/src/poly.rs:56-64,133-139) and the direct
 * evaluate_over_domain_by_ref / interpolate calls at crates/plonk/src/plonk/protocol.rs:82-106,140-141,158-159,195.
 * data: b vectors of 2^log_n elements of `field`, transformed in place, natural order in and out; the inverse
 * includes the factor N^-1.  coset_offset (4 words, Montgomery) evaluates on g*<omega> / interpolates from it
 * (API addition: the reference has no coset call site); NULL = the plain subgroup domain.
 * omega_N = (5^((p-1)/2^32))^(2^(32-log_n)) as in ark-ff (pinned by ivc/mod.rs:56,113). */
int halo_ntt(int field, uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null);
int halo_ntt_dev(int field, void* d_data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null);
/* evals[i] = p(omega^i) for a coefficient vector of ANY length: zero-padded when len < N, reduced modulo X^N - 1
 * when len > N, as ark-poly's evaluate_over_domain_by_ref does (protocol.rs:140-141 relies on the latter). */
int halo_poly_evaluate(int field, const uint64_t* coeffs, size_t len, unsigned log_n, uint64_t* out_evals);

/* coefficients of a*b (la + lb - 1 of them)  ≙ `&a * &b` on DensePolynomial (protocol.rs:132-139,198-199; pcdl.rs:215):
 * two forward transforms, a pointwise product and one inverse transform on the domain of size next_pow2(la+lb-1) */
int halo_poly_mul(int field, const uint64_t* a, size_t la, const uint64_t* b, size_t lb, uint64_t* out);
/* ≙ scalar_dot (group.rs:43-45) and construct_powers (group.rs:58-66): out[i] = z^i, i < n */
int halo_scalar_dot(int field, const uint64_t* xs, const uint64_t* ys, size_t n, uint64_t out[4]);
int halo_construct_powers(int field, const uint64_t z[4], size_t n, uint64_t* out);
/* No counterpart in the reference: added for callers that cannot allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as halo_scalar_dot --
 * out = sum_{i<n} xs[i] * ys[i] in `field` -- with every pointer a DEVICE pointer (d_out: one element, 32 bytes) and
 * nothing allocated: the per-block partial sums live in a fixed 128 KiB file-scope device array of the library (4096
 * slots, one per 4096 elements) and the sum is finished on the device.  Two kernel launches on the library's stream
 * (halo_set_stream); the call returns without synchronising -- order later work on that stream or call
 * halo_device_synchronize.  The first call on a device looks the array's address up (cudaGetSymbolAddress), so make
 * it before starting a graph capture.
 * Limit: n <= 2^24 elements per call; larger n -> HALO_EINVAL (use halo_scalar_dot).
 * Every call shares the one scratch array, which keeps the previous call's contents (a call reads only the slots it
 * has written itself, so nothing needs clearing): calls on one stream are ordered, calls that OVERLAP on the GPU
 * (several streams, concurrent graph launches) race. */
int halo_scalar_dot_noalloc_dev(int field, const void* d_xs, const void* d_ys, size_t n, void* d_out);

/* ---- PLONK prover / verifier -------------------------------------------------------------------------------
 * ≙ PlonkProof::naive_prover (crates/plonk/src/plonk/protocol.rs:64-355) on what Trace::consume produces
 * (crates/plonk/src/circuit/trace.rs:234-283): `polys` = the coefficient vectors witness.polys.{qs[10], ws[16],
 * rs[15], ids[8], sigmas[8]} in that order, each zero-padded to n = 2^lg_n elements (57 x n x 4 words);
 * `w_evals` = witness.w_evals[0..3] (3 x n x 4 words, row i at omega^(i+1) as Evals stores them);
 * `public_inputs` = public_inputs.public_inputs; `acc_prev` = public_inputs.acc_prev as a flat instance.
 * No hiding is used on this path (protocol.rs:295-306 pass None), so no randomness is consumed.
 * Proof layout (halo_plonk_proof_words(lg_n) = 828 + 72 lg_n words):
 *   vs  = ws[16] rs[15] qs[10] ts[16] ids[8] sigmas[8] z z_omega w_omegas[3]        (4 words each)
 *   Cs  = ws[16] ts[16] z                                                           (12 words each, Jacobian)
 *   pis = r, r_omega   (EvalProof: Ls[12L] Rs[12L] U[12] c[4] hiding[1] C_bar[12] w_prime[4])
 *   acc_next           (flat instance)
 * Errors: HALO_EINVAL where the reference panics (a zero denominator in the grand product), HALO_EPARAMS without
 * public parameters of at least n generators. */
size_t halo_plonk_proof_words(unsigned lg_n);
int halo_plonk_prove(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                     size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof);
/* The gate constraints on H: out_evals[j] = f_GC(omega^j), j < n (n x 4 words), i.e. the F_GC of
 * crates/plonk/src/plonk/protocol.rs:170-241 evaluated on the rows, without the permutation terms.  All zero exactly when
 * the witness satisfies every gate row (the quotient then takes the 8-coset path).  Same inputs as halo_plonk_prove; a
 * set-up / diagnostic call (the synthetic all-gates circuit of the benchmark solves its constant selector with it). */
int halo_plonk_gate_residual(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                             size_t n_public, uint64_t* out_evals);
/* The circuit's own columns kept on the device across proofs of the same circuit -- the analogue of the reference
 * holding `x: PlonkCircuit` and the Trace's polynomials for every step of an IVC chain (crates/plonk/src/frontend/ivc/
 * mod.rs:640-660 proves one circuit per curve over and over).  halo_plonk_circuit_load takes the same 57-column
 * `polys` as halo_plonk_prove and keeps qs[10], rs[15], ids[8], sigmas[8] resident (41 x n x 32 B); *handle >= 1.
 * halo_plonk_prove_circuit then needs only the 16 witness coefficient columns (`w_polys`, 16 x n x 4 words) and
 * produces the identical proof.  Handles are released with halo_plonk_circuit_free or by halo_shutdown. */
int halo_plonk_circuit_load(int curve, unsigned lg_n, const uint64_t* polys, int* handle);
int halo_plonk_circuit_free(int handle);
int halo_plonk_prove_circuit(int handle, const uint64_t* w_polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                             size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof);
/* ≙ verify_succinct (protocol.rs:357-491) when succinct_only != 0, else verify (:493-502 = succinct + acc::decider).
 * circuit_C_qs = circuit.Cs.qs (10 x 12 words); public_input_count = circuit.public_input_count.
 * HALO_OK accept, HALO_EREJECT where the reference's ensure! fails. */
int halo_plonk_verify(int curve, unsigned lg_n, const uint64_t* circuit_C_qs, const uint64_t* public_inputs, size_t n_public,
                      size_t public_input_count, const uint64_t* acc_prev, const uint64_t* proof, int succinct_only);

/* ---- benchmarks of the arithmetic itself (denominator of the integer roofline) -------------------------- */
/* Runs `iters` dependent Montgomery multiplications in each of blocks*threads threads (4 independent chains per
 * thread); returns elapsed milliseconds measured with CUDA events in *ms and a checksum in out[4]. */
int halo_bench_modmul(int field, int blocks, int threads, int iters, float* ms, uint64_t out[4]);
/* raw IMAD.WIDE.U32 issue rate: `iters` x 8 independent wide multiply-adds per thread */
int halo_bench_imad(int blocks, int threads, int iters, float* ms, uint64_t* out);
/* issue rate of other candidate pipes, `iters` x 8 independent operations per thread: mode 1 = DFMA,
 * 2 = DFMA and IMAD.WIDE interleaved 1:1 (8 + 8 per iteration), 3 = 32-bit IMAD, 4 = carry-chained IADD */
int halo_bench_pipe(int mode, int blocks, int threads, int iters, float* ms, uint64_t* out);

/* ---- instrumentation ------------------------------------------------------------------------------------- */
/* When enabled, every MSM records CUDA events between its pipeline stages on the stream it launches on;
 * halo_profile_read returns the number of stages of the LAST call and fills their durations (ms) and names. */
int halo_profile_enable(int on);
int halo_profile_read(float* ms, const char** names, int max_stages);
/* number of kernels this library has launched since it was loaded */
unsigned long long halo_launch_count(void);

/* ---- synthetic inputs for benchmarks and property tests ---------------------------------------------------- */
/* out[i] = m_i * G (affine, Montgomery) for i in [first, first+n), m_i = halo_gen_point_multiplier(seed, i), a
 * 64-bit odd integer; G = (-1, 2) is the generator of both Pasta curves.  Because the discrete logs are known,
 * sum_i s_i * P_i == (sum_i s_i * m_i) * G gives a size-independent check of an MSM result. */
uint64_t halo_gen_point_multiplier(uint64_t seed, uint64_t i);
int halo_gen_points_dev(int curve, uint64_t seed, size_t first, size_t n, void* d_out_xy);
int halo_gen_points(int curve, uint64_t seed, size_t first, size_t n, uint64_t* out_xy);

/* ---- elementwise helpers used by tests -------------------------------------------------------------------- */
/* op: 0 mul, 1 add, 2 sub, 3 inverse(a), 4 to_canonical(a), 5 from_canonical(a); n elements, on the GPU */
int halo_field_op(int field, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* HALO_B200_H */
