/*
 * pcseg_ops.h — C ABI of libpcseg_ops.so (sm_100a), the drop-in boundary for the tf_ops/
 * neighbor-operator layer of liuyuan-pal/PointCloudSegmentation.
 *
 * Every entry point replaces ONE host launcher of the reference (the function a TF OpKernel::Compute
 * calls); the citation after "replaces:" is the reference file:line.  Conventions:
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - feature tensors are row-major float32, index tensors int32 (the reference reinterprets int32
 *     as unsigned, TFNeighborScatter.cc:66-68; values are non-negative either way);
 *   - the library never allocates, frees or synchronises: outputs and workspaces are the caller's,
 *     all work is enqueued on `stream` (a cudaStream_t passed as void*);
 *   - return value 0 = success; non-zero = error, text via pcs_last_error() (thread-local).
 *     The reference prints and exit()s the process on CUDA errors (TFCudaCommon.h:14-22); we never do;
 *   - inputs are never written.
 *   - CSR vocabulary: pn centres, csum edges, nidxs[csum] neighbour per edge, nidxs_lens[pn] degree,
 *     nidxs_bgs[pn] exclusive prefix sum of lens, cidxs[csum] centre per edge.
 */
#ifndef PCSEG_OPS_H
#define PCSEG_OPS_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void *pcs_stream_t; /* cudaStream_t */

/* ---- library ---------------------------------------------------------------------------- */
const char *pcs_last_error(void);
int pcs_version(void);
/* number of kernels this library has launched since load (for bench.py's gpu_launches claim) */
int64_t pcs_launch_count(void);

/* ---- a1: NeighborScatter.  replaces: neighborScatterGPU TFNeighborKernel.cu:255-288
 * sfeats[bg_i+j,:] = ifeats[nidxs[bg_i+j],:] (- ifeats[i,:] if use_diff)                     */
int pcs_neighbor_scatter(const float *ifeats, const int32_t *nidxs, const int32_t *nidxs_lens,
                         const int32_t *nidxs_bgs, const int32_t *cidxs /* nullable: centre per edge */,
                         int64_t pn, int64_t ifn, int64_t csum, int use_diff, float *sfeats,
                         pcs_stream_t stream);

/* ---- helper: cidxs[bg_i+j] = i, the COO row the reference's searches also return as `cens`
 * (computeNeighborIdxsKernel PoolingIndex/SearchNeighborhood.cu:57-88).  Lets a1 with use_diff run
 * edge-parallel; pass NULL to a1 to walk rows per centre instead.                              */
int pcs_csr_cidxs(const int32_t *nidxs_lens, const int32_t *nidxs_bgs, int64_t pn, int64_t csum,
                  int32_t *cidxs, pcs_stream_t stream);

/* ---- transpose CSR (edges grouped by neighbour, ascending edge id) — what makes a2 deterministic.
 * No reference counterpart (the reference uses atomicAdd, TFNeighborKernel.cu:83,112,121).
 * tperm[csum]: edge ids sorted (stably) by nidxs; tbgs[pn+1]: row starts in tperm.            */
size_t pcs_csr_transpose_workspace_bytes(int64_t pn, int64_t csum);
int pcs_csr_transpose(const int32_t *nidxs, int64_t pn, int64_t csum, int32_t *tperm, int32_t *tbgs,
                      void *workspace, size_t workspace_bytes, pcs_stream_t stream);


/* Same result for a SYMMETRIC CSR with ascending rows (the output of the radius / annulus searches a18, a19:
 * j in N(i) <=> i in N(j)): tbgs = bgs and one binary search per edge instead of a radix sort of all edges.
 * cidxs: centre per edge.  *fail_flag (device int32, zeroed by the caller) is set to 1 when an edge has no
 * mirror, i.e. the graph was not symmetric and tperm is invalid.                                */
int pcs_csr_transpose_symmetric(const int32_t *nidxs, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                                const int32_t *cidxs, int64_t pn, int64_t csum, int32_t *tperm, int32_t *tbgs,
                                int32_t *fail_flag, pcs_stream_t stream);


/* ---- a2: NeighborGather.  replaces: neighborGatherGPU TFNeighborKernel.cu:292-332
 * ifeats[n,:] = sum_{e: nidxs[e]==n} sfeats[e,:]  (- sum_{e in seg(n)} sfeats[e,:] if use_diff)
 * summed in ascending edge order (deterministic).                                             */
int pcs_neighbor_gather(const float *sfeats, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                        const int32_t *tperm, const int32_t *tbgs, int64_t pn, int64_t ifn,
                        int64_t csum, int use_diff, float *ifeats, pcs_stream_t stream);

/* ---- a3: NeighborSumFeatGather.  replaces: neighborSumFeatGatherGPU (declared
 * TFNeighborKernel.h:84-93, kernel TFNeighborKernelNew.cu:4-27, launcher missing upstream)
 * ogfeats[i,:] = sum_{j<len_i} ifeats[bg_i+j,:]                                               */
int pcs_neighbor_sum_feat_gather(const float *ifeats, const int32_t *nidxs_lens,
                                 const int32_t *nidxs_bgs, int64_t pn, int64_t fd, int64_t csum,
                                 float *ogfeats, pcs_stream_t stream);

/* ---- a4: NeighborSumFeatScatter.  replaces: neighborSumFeatScatterGPU TFNeighborKernelNew.cu:50-79
 * osfeats[e,:] = igfeats[cidxs[e],:].  nidxs_lens / nidxs_bgs (the op's other two inputs,
 * TFNeighborSumFeatScatter.cc:12-20; nullable here) must describe the same CSR as cidxs with
 * nidxs_bgs = exclusive scan of nidxs_lens; when given, rows >= 8 floats wide are written as a
 * per-centre row broadcast instead of a per-edge gather.                                          */
int pcs_neighbor_sum_feat_scatter(const float *igfeats, const int32_t *cidxs, const int32_t *nidxs_lens,
                                  const int32_t *nidxs_bgs, int64_t pn, int64_t fd, int64_t csum,
                                  float *osfeats, pcs_stream_t stream);

/* ---- a5: NeighborMaxFeatGather.  replaces: neighborMaxFeatGatherGPU TFNeighborKernelNew.cu:141-172
 * per-voxel channel max and the offset (within the segment) of the FIRST maximum (strict <).
 * Empty segment: gfeats = 0 (reference memset) and max_idxs = 0 (reference leaves garbage).    */
int pcs_neighbor_max_feat_gather(const float *ifeats, const int32_t *vlens, const int32_t *vbegs,
                                 int64_t pn1, int64_t pn2, int64_t fd, float *gfeats,
                                 int32_t *max_idxs, pcs_stream_t stream);

/* ---- a6: NeighborMaxFeatScatter.  replaces: neighborMaxFeatScatterGPU TFNeighborKernelNew.cu:175-207
 * sfeats = 0; sfeats[vbegs[i]+max_idxs[i,f], f] = gfeats[i,f].
 * Precondition (the CSR invariant every caller satisfies): vbegs = exclusive scan of vlens and
 * max_idxs[i,f] < vlens[i].  Rows >= sum(vlens) are zero-filled.                               */
int pcs_neighbor_max_feat_scatter(const float *gfeats, const int32_t *max_idxs, const int32_t *vbegs,
                                  const int32_t *vlens, int64_t pn1, int64_t pn2, int64_t fd,
                                  float *sfeats, pcs_stream_t stream);

/* ---- a7: LocationWeightSum.  replaces: locWSumForwardGPU TFNeighborKernel.cu:415-434        */
int pcs_location_weight_sum(const float *lw, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                            int64_t pn, int64_t m, int64_t csum, float *lw_sum, pcs_stream_t stream);

/* ---- a8: LocationWeightSumBackward.  replaces: locWSumBackwardGPU TFNeighborKernel.cu:436-454
 * dlw[bg_i+j,:] = dlw_sum[i,:]; rows of dlw outside every segment are zero (reference memset).  */
int pcs_location_weight_sum_backward(const float *dlw_sum, const int32_t *nidxs_lens,
                                     const int32_t *nidxs_bgs, int64_t pn, int64_t m, int64_t csum,
                                     float *dlw, pcs_stream_t stream);

/* ---- a9: LocationWeightFeatSum.  replaces: locWFeatSumForwardGPU TFNeighborKernel.cu:334-370
 * otfeats_sum[i,a,o] = sum_j lw[bg_i+j,a] * tfeats[bg_i+j,a,o]                                */
int pcs_location_weight_feat_sum(const float *lw, const float *tfeats, const int32_t *nidxs_lens,
                                 const int32_t *nidxs_bgs, int64_t pn, int64_t m, int64_t ofn,
                                 int64_t csum, float *otfeats_sum, pcs_stream_t stream);

/* ---- a10: LocationWeightFeatSumBackward.  replaces: locWFeatSumBackwardGPU TFNeighborKernel.cu:372-412
 * dtfeats[e,a,o] = dsum[i,a,o]*lw[e,a];  dlw[e,a] = sum_o dsum[i,a,o]*tfeats[e,a,o]  (no atomics) */
int pcs_location_weight_feat_sum_backward(const float *lw, const float *tfeats, const float *dsum,
                                          const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                                          int64_t pn, int64_t m, int64_t ofn, int64_t csum,
                                          float *dlw, float *dtfeats, pcs_stream_t stream);

/* ---- a11: EliminateCenter.  replaces: eliminateCenterGPU TFNeighborKernelNew.cu:372-392
 * Drops entries equal to the row's own centre; olens = lens-1, obgs = bgs-i.  Precondition (as
 * upstream, TFNeighborEliminateCenter.cc:62): every row holds its centre exactly once.
 * At most lens-1 entries are written per row (upstream would overrun the next row instead).     */
int pcs_eliminate_center(const int32_t *inidxs, const int32_t *inidxs_lens, const int32_t *inidxs_bgs,
                         int64_t pn, int64_t csum, int32_t *onidxs, int32_t *onidxs_lens,
                         int32_t *onidxs_bgs, int32_t *ocidxs, pcs_stream_t stream);

/* ---- a12: ComputeVoxelIndex.  replaces: computeVoxelIdxImpl PoolingIndex/ComputeVoxelIdx.cu:29-49
 * voxel_idxs[i,k] = (uint32) floor((pts[i,k]-min_k)/voxel_len), IEEE fp32 divide, saturating.
 * min_* are passed already formed in fp32 by the caller (ComputeVoxelIdx.cc:66-67).             */
int pcs_compute_voxel_index(const float *pts, int64_t pn, int64_t pt_stride, float voxel_len,
                            float min_x, float min_y, float min_z, int32_t *voxel_idxs,
                            pcs_stream_t stream);

/* ---- a13/a14: ComputePermutationInfo[WithClass].  replaces: computePermutationInfoImpl
 * PoolingIndex/ComputePermutationInfo.cu:91-155 and ...WithClassesImpl :157-229.
 * Stable sort by key (x<<32)+(y<<16)+z  [ ((x<<48)+(y<<32)+(z<<16)+class) with classes ].
 * block_ids (nullable, extension for block-diagonal batches): sorted as the most significant
 * field, so voxels never merge across blocks; must be non-decreasing-agnostic (any order).
 * lens/begs are caller buffers of pn entries (upper bound); only the first *vn are defined.
 * vn_dev: device int32[2]: [0] receives the voxel count, [1] status flags (caller copies both to
 * the host in one transfer): bit0 = some key needs more than the 48 bits that were sorted (voxel
 * coordinate outside [0,65535]) -> call pcs_compute_permutation_info64, which sorts the
 * full signed 64-bit key like thrust does; bit1 = coordinate out of range while block_ids given.   */
size_t pcs_compute_permutation_info_workspace_bytes(int64_t pn);
int pcs_compute_permutation_info(const int32_t *voxel_idxs, const int32_t *classes,
                                 const int32_t *block_ids, int64_t pn, int32_t *o2p, int32_t *p2o,
                                 int32_t *lens, int32_t *begs, int32_t *cens, int32_t *vn_dev,
                                 void *workspace, size_t workspace_bytes, pcs_stream_t stream);
int pcs_compute_permutation_info64(const int32_t *voxel_idxs, const int32_t *classes,
                                   const int32_t *block_ids, int64_t pn, int32_t *o2p, int32_t *p2o,
                                   int32_t *lens, int32_t *begs, int32_t *cens, int32_t *vn_dev,
                                   void *workspace, size_t workspace_bytes, pcs_stream_t stream);

/* ---- a15: PermutateFeature (+Backward: same kernel with the inverse permutation).
 * replaces: permutateFeature<T> PoolingIndex/PermutateFeature.cu:21-50.  T is any 4-byte type.
 * out[i,:] = feats[idxs[i],:]                                                                  */
int pcs_permutate_feature(const void *feats, const int32_t *idxs, int64_t pn, int64_t ps, void *out,
                          pcs_stream_t stream);

/* ---- a16: ComputeDiffXyz.  replaces: computeDiffXYZImpl PoolingIndex/ComputeDiffXYZ.cu:22-38  */
int pcs_compute_diff_xyz(const float *xyzs, const float *cxyzs, const int32_t *cidxs, int64_t pn1,
                         int64_t pn2, float *dxyzs, pcs_stream_t stream);

/* ---- a17: ComputeRepermutationInfo.  replaces: computeRepermutationInfoImpl
 * PoolingIndex/ComputeRepermutationInfo.cu:52-88                                                */
size_t pcs_compute_repermutation_info_workspace_bytes(int64_t pn2);
int pcs_compute_repermutation_info(const int32_t *o2p_idxs2, const int32_t *lens, const int32_t *begs,
                                   int64_t pn1, int64_t pn2, int32_t *reper_o2p1, int32_t *reper_p2o1,
                                   int32_t *reper_lens, int32_t *reper_begs, int32_t *reper_cens,
                                   void *workspace, size_t workspace_bytes, pcs_stream_t stream);

/* ---- a18/a19: SearchNeighborhoodBruteForce[Range].  replaces: searchNeighborhoodCount[Range]Impl
 * + searchNeighborhood[Range]Impl PoolingIndex/SearchNeighborhood.cu:126-223.
 * Hit iff d2 < sq_max (&& d2 > sq_min when use_min), d2 = fma(dz,dz,fma(dx,dx,dy*dy)) — the
 * contraction nvcc emits for SearchNeighborhood.cu:24.  Rows list hits in ascending index.
 * Two phases like SearchNeighborhood.cc:76-83: count (lens, begs, total on device), then fill.
 * block_begs (nullable, extension): [nb+1] point offsets of independent blocks; points only see
 * candidates of their own block (indices stay global).  max_block_pn = largest block size.
 * Every call needs a workspace of pcs_search_workspace_bytes(pn) (per-32-point bounding boxes used
 * to skip far-away candidates conservatively: the output equals plain brute force bit for bit).  */
size_t pcs_search_workspace_bytes(int64_t pn);
int pcs_search_neighborhood_count(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                  int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                  int32_t *lens, int32_t *begs, int32_t *total_dev, void *workspace,
                                  size_t workspace_bytes, pcs_stream_t stream);
int pcs_search_neighborhood_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                 int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                 const int32_t *begs, int32_t *idxs, int32_t *cens, void *workspace,
                                 size_t workspace_bytes, pcs_stream_t stream);

/* ---- a20: SearchNeighborhoodFixedBruteForce[Range].  replaces: searchNeighborhoodFixed[Range]Impl
 * + computeNeighborhoodStats PoolingIndex/SearchNeighborhoodFixed.cu:105-152.
 * First `fixed_size` hits in index order, padded with the last hit (or the centre if none).     */
int pcs_search_neighborhood_fixed(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                  int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                  int64_t fixed_size, int32_t *idxs, int32_t *lens, int32_t *begs,
                                  int32_t *cens, void *workspace, size_t workspace_bytes,
                                  pcs_stream_t stream);

/* ---- extension (fused edge-conv): row-wise dense-concat MLP of the reference's
 * pointnet_conv (model_pointnet.py:10-24) after distributing the linear layers over the concat:
 *   u = [u1(A)|u2(B)|u3(C)|uo(O)] per row;  h1=relu(u1); h2=relu(u2+W2 h1); h3=relu(u3+W3[h2,h1]);
 *   y = uo + Wo[h3,h2,h1].   w2 [B,A], w3 [C,B+A], wo [O,C+B+A] row-major.
 * backward: du [rows,A+B+C+O] = [dpre1|dpre2|dpre3|dy], hcat [rows,A+B+C] = [h3|h2|h1] (recomputed).
 * Supported widths: pcs_dense_mlp3_supported(A,B,C,O) != 0.  No reference CUDA counterpart.        */
int pcs_dense_mlp3_supported(int64_t A, int64_t B, int64_t C, int64_t O);
int pcs_dense_mlp3_forward(const float *u, const float *w2, const float *w3, const float *wo, int64_t rows,
                           int64_t A, int64_t B, int64_t C, int64_t O, float *y, pcs_stream_t stream);
int pcs_dense_mlp3_backward(const float *u, const float *dy, const float *w2, const float *w3,
                            const float *wo, int64_t rows, int64_t A, int64_t B, int64_t C, int64_t O,
                            float *du, float *hcat, pcs_stream_t stream);

/* ---- tuning knobs (dispatch thresholds; defaults are the measured best on B200) ------------------------
 * "scatter_walk_min_c_diff" / "_diff_odd" / "_copy" / "_copy_odd": pcs_neighbor_scatter uses the
 * centre-walking kernel for rows at least this wide (C % 4 == 0 / other widths; use_diff / copy);
 * a negative value means never.  Results are identical either way.
 * "mlp_tensor_cores": 1 (default) runs pcs_dense_mlp3_* on the tensor cores with error-compensated
 * TF32 (three MMAs per product, ~1e-6 relative), 0 on the fp32 FFMA pipe.                           */
int pcs_set_option(const char *name, int64_t value);

/* ---- bf16-storage twins of the feature movers (a1-a10, a15).  The reference instantiates its launcher templates
 * for FLT = float only (e.g. TFNeighborKernel.cu:456-467); these are the same launchers for rows stored as
 * bfloat16 (void* = __nv_bfloat16*): 16-byte lanes carry 8 columns, every sum / product accumulates in fp32 and is
 * rounded to nearest-even once on store, copies and max / arg-max are exact (bf16 -> fp32 is a 16-bit shift), so
 * index outputs equal the fp32 entry's on the same (bf16-representable) inputs.  Same argument meaning as the
 * fp32 entry of the same name.                                                                             */
int pcs_neighbor_scatter_bf16(const void *ifeats, const int32_t *nidxs, const int32_t *nidxs_lens,
                              const int32_t *nidxs_bgs, const int32_t *cidxs, int64_t pn, int64_t ifn, int64_t csum,
                              int use_diff, void *sfeats, pcs_stream_t stream);
int pcs_neighbor_gather_bf16(const void *sfeats, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                             const int32_t *tperm, const int32_t *tbgs, int64_t pn, int64_t ifn, int64_t csum,
                             int use_diff, void *ifeats, pcs_stream_t stream);
int pcs_neighbor_sum_feat_gather_bf16(const void *ifeats, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                                      int64_t pn, int64_t fd, int64_t csum, void *ogfeats, pcs_stream_t stream);
int pcs_neighbor_sum_feat_scatter_bf16(const void *igfeats, const int32_t *cidxs, const int32_t *nidxs_lens,
                                       const int32_t *nidxs_bgs, int64_t pn, int64_t fd, int64_t csum,
                                       void *osfeats, pcs_stream_t stream);
int pcs_neighbor_max_feat_gather_bf16(const void *ifeats, const int32_t *vlens, const int32_t *vbegs, int64_t pn1,
                                      int64_t pn2, int64_t fd, void *gfeats, int32_t *max_idxs, pcs_stream_t stream);
int pcs_neighbor_max_feat_scatter_bf16(const void *gfeats, const int32_t *max_idxs, const int32_t *vbegs,
                                       const int32_t *vlens, int64_t pn1, int64_t pn2, int64_t fd, void *sfeats,
                                       pcs_stream_t stream);
int pcs_location_weight_sum_bf16(const void *lw, const int32_t *nidxs_lens, const int32_t *nidxs_bgs, int64_t pn,
                                 int64_t m, int64_t csum, void *lw_sum, pcs_stream_t stream);
int pcs_location_weight_sum_backward_bf16(const void *dlw_sum, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                                          int64_t pn, int64_t m, int64_t csum, void *dlw, pcs_stream_t stream);
int pcs_location_weight_feat_sum_bf16(const void *lw, const void *tfeats, const int32_t *nidxs_lens,
                                      const int32_t *nidxs_bgs, int64_t pn, int64_t m, int64_t ofn, int64_t csum,
                                      void *otfeats_sum, pcs_stream_t stream);
int pcs_location_weight_feat_sum_backward_bf16(const void *lw, const void *tfeats, const void *dtfeats_sum,
                                               const int32_t *nidxs_lens, const int32_t *nidxs_bgs, int64_t pn,
                                               int64_t m, int64_t ofn, int64_t csum, void *dlw, void *dtfeats,
                                               pcs_stream_t stream);
int pcs_permutate_feature_bf16(const void *feats, const int32_t *idxs, int64_t pn, int64_t ps, void *out,
                               pcs_stream_t stream);

/* ---- extension: the edge MLP feeding the segmented max (pointnet_conv = MLP then graph_pool) --------
 * pcs_dense_mlp3_backward_pooled: pcs_dense_mlp3_backward with dy given implicitly by the pooled
 *   gradient: dy[r,f] = dpool[i,f] if max_idxs[i,f] == r - nidxs_bgs[i] else 0, i = cidxs[r]
 *   (what a6 NeighborMaxFeatScatter would materialise).  dpool [pn,O] f32, max_idxs [pn,O] i32.
 *   wpart (nullable): [P, C*(B+A) + B*A] with P = pcs_dense_mlp3_wgrad_partials(rows,A,B,C,O) > 0; the
 *   sum over P is [dW3 (row-major [C,B+A]) | dW2 ([B,A])] = dpre3^T [h2|h1], dpre2^T h1, accumulated
 *   inside the same kernel.  P == 0 means unavailable (mlp_tensor_cores = 0): pass NULL.
 * pcs_pool_linear_wgrad: partial [P,O,J], P = pcs_pool_linear_wgrad_partials(pn,O); the sum over P is
 *   dWo[f,:] = sum_i dpool[i,f] * x[nidxs_bgs[i] + max_idxs[i,f], :]   (x [csum,J], J in {32,48,64}),
 *   i.e. dy^T x without touching the rows that lost the max.  Empty neighbourhoods contribute 0.   */
int pcs_dense_mlp3_backward_pooled(const float *u, const float *dpool, const int32_t *max_idxs,
                                   const int32_t *cidxs, const int32_t *nidxs_bgs, const float *w2,
                                   const float *w3, const float *wo, int64_t rows, int64_t A, int64_t B,
                                   int64_t C, int64_t O, float *du, float *hcat, float *wpart,
                                   pcs_stream_t stream);
int64_t pcs_dense_mlp3_wgrad_partials(int64_t rows, int64_t A, int64_t B, int64_t C, int64_t O);
int64_t pcs_pool_linear_wgrad_partials(int64_t pn, int64_t O);
int pcs_pool_linear_wgrad(const float *dpool, const int32_t *max_idxs, const int32_t *nidxs_lens,
                          const int32_t *nidxs_bgs, const float *x, int64_t pn, int64_t csum, int64_t O,
                          int64_t J, float *partial, pcs_stream_t stream);

/* ---- extension: column sums of a tall matrix (csrc/column_sum.cu) -----------------------------------
 * out[c] = sum_r x[r,c], x [rows,C] row-major f32: the bias gradient of the point-level linear layers
 * (tf.nn.bias_add's gradient in the reference's stock-TF glue) and the fold of per-CTA partials.
 * partial: scratch [P,C] with P = pcs_column_sum_partials(rows,C); P == 0 means a single pass (pass NULL).
 * Fixed summation order (no atomics).  rows == 0 writes zeros.                                       */
int64_t pcs_column_sum_partials(int64_t rows, int64_t C);
int pcs_column_sum(const float *x, int64_t rows, int64_t C, float *partial, float *out, pcs_stream_t stream);

/* ---- extension: edge-tensor assembly of the split edge convolution (csrc/edge_conv.cu) -------------
 * What model_pointnet.py:10-24 builds with graph_concat_scatter + concat + fully_connected, after
 * distributing every layer's weight over the concat [feats[centre] | feats[neighbour] | sxyz]:
 *   u[e,:] = pa[cidxs[e],:] + pb[nidxs[e],:] + wx * sxyz[e,:]      pa, pb [pn,H]; wx [H,3]; u [csum,H]
 * pa and pb have row stride ldp floats (ldp = 2H when they are the two halves of one [pn,2H] product).
 * H % 4 == 0, H <= 256, ldp % 4 == 0; pa, pb, u 16-byte aligned.  With (nidxs_lens, nidxs_bgs) the
 * centre-walking kernel runs (cidxs may then be NULL); without them the edge-parallel one (needs cidxs).
 * pcs_edge_xyz_wgrad: partial [P,H,3], P = pcs_edge_xyz_wgrad_partials(csum,H); summing partial over
 * P (fixed order) gives d wx = du^T sxyz.  No reference CUDA counterpart.                          */
int pcs_edge_combine_forward(const float *pa, const float *pb, const float *sxyz, const float *wx,
                             const int32_t *cidxs, const int32_t *nidxs, const int32_t *nidxs_lens,
                             const int32_t *nidxs_bgs, int64_t pn, int64_t csum, int64_t H, int64_t ldp, float *u,
                             pcs_stream_t stream);
int64_t pcs_edge_xyz_wgrad_partials(int64_t csum, int64_t H);
int pcs_edge_xyz_wgrad(const float *du, const float *sxyz, int64_t csum, int64_t H, float *partial,
                       pcs_stream_t stream);

/* ---- extension: point-level linear layer on the tcgen05 tensor cores (csrc/linear_tc.cu) ------------
 * y[M,N] = x[M,K] . B^T (+ bias[N]) (+ addend[M,N]) (relu if act == 1),  B[n,k] = w[n*w_stride_n + k*w_stride_k]
 * (so a [N,K] weight, its transpose or a column slice of a wider weight are all passed as they are).
 * x / y / addend row-major fp32 with row strides ldx / ldy / ldadd (elements); any K and N.
 * Stands where the reference calls tf.contrib.layers.fully_connected between the custom ops
 * (model_pointnet.py:10-24,56-69, model.py:762-799); no reference CUDA counterpart (stock TF GEMM).
 * fp32 accuracy: every product is three TF32 MMAs on an error-compensated hi/lo split, fp32 accumulation
 * in TMEM.  workspace: pcs_linear_tc_workspace_bytes(N,K) bytes, 16-byte aligned (the split weight image). */
size_t pcs_linear_tc_workspace_bytes(int64_t N, int64_t K);
int pcs_linear_tc(const float *x, int64_t M, int64_t K, int64_t ldx, const float *w, int64_t N,
                  int64_t w_stride_n, int64_t w_stride_k, const float *bias /* nullable */,
                  const float *addend /* nullable */, int64_t ldadd, int act, float *y, int64_t ldy,
                  void *workspace, size_t workspace_bytes, pcs_stream_t stream);

/* Weight (and bias) gradient of the same layer (csrc/linear_tc_wgrad.cu): dW[N,K] = g[M,N]^T . x[M,K], and with
 * with_bias != 0 one more column db[n] = sum_m g[m,n] (x is extended by a column of ones on the fly).
 * partial: [S][N][K + with_bias] floats, S = pcs_linear_tc_wgrad_partials(M,N,K,with_bias) row splits; the caller
 * folds them over S in fixed order (pcs_column_sum): deterministic, no atomics.  Same 3xTF32 arithmetic.      */
int64_t pcs_linear_tc_wgrad_partials(int64_t M, int64_t N, int64_t K, int with_bias);
int pcs_linear_tc_wgrad(const float *g, int64_t M, int64_t N, int64_t ldg, const float *x, int64_t K, int64_t ldx,
                        int with_bias, float *partial, pcs_stream_t stream);

/* ---- extension: classifier glue (csrc/relu_concat_dropout.cu), classifier_v3 model.py:783-793 ------------
 * out[M, Na+Nb] (row-major, contiguous) = dropout(concat([relu(a[M,Na]), b[M,Nb]], 1), keep_prob) in one pass;
 * keep_prob >= 1 disables dropout (inference).  Kept values are scaled by 1/keep_prob (tf.nn.dropout).  The
 * keep bits are a pure function of (seed, element index): Philox4x32-10, e = row*(Na+Nb) + col, counter
 * (32*(e/128) + e%32, 0), key = seed, word (e%128)/32, keep iff word < keep_prob*2^32 — the backward regenerates them, no mask is stored.
 * backward: da[M,Na] = g * (out > 0 ? 1/keep_prob : 0), db[M,Nb] (nullable) = g * keep / keep_prob.          */
int pcs_relu_concat_dropout_forward(const float *a, int64_t M, int64_t Na, int64_t lda, const float *b /* nullable if Nb == 0 */,
                                    int64_t Nb, int64_t ldb, float keep_prob, uint64_t seed, float *out,
                                    pcs_stream_t stream);
int pcs_relu_concat_dropout_backward(const float *g, const float *out, int64_t M, int64_t Na, int64_t Nb, float keep_prob,
                                     uint64_t seed, float *da, int64_t ldda, float *db /* nullable */, int64_t lddb,
                                     pcs_stream_t stream);

/* ---- several radius / annulus graphs over the same points from one traversal (extension; each graph's output is
 * what pcs_search_neighborhood_count/fill give for it: SearchNeighborhood.cu:5-123, SearchNeighborhood.cc:30-169).
 * 1 <= ng <= 4; sq_min / sq_max / use_min are HOST arrays of ng entries; lens, begs: [ng, pn] graph-major;
 * totals_dev: ng device int32 edge totals (one D2H read sizes all graphs); idxs, cens: HOST arrays of ng device
 * pointers.  workspace >= pcs_search_workspace_bytes(pn).                                                        */
int pcs_search_multi_count(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                           int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min, int32_t *lens,
                           int32_t *begs, int32_t *totals_dev, void *workspace, size_t workspace_bytes, pcs_stream_t stream);
int pcs_search_multi_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                          int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min, const int32_t *begs,
                          int32_t *const *idxs, int32_t *const *cens, void *workspace, size_t workspace_bytes,
                          pcs_stream_t stream);
/* The same two steps with ONE candidate traversal instead of two: the count pass also records each centre's first maxk
 * hits of every graph in pad [ng, pn, maxk] (ascending, as the CSR rows); the fill step copies the rows that fit
 * (lens <= maxk) from pad and only the warps owning a longer row traverse again.  Outputs identical to count / fill.  */
int pcs_search_multi_count_pad(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                               int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min, int32_t *lens,
                               int32_t *begs, int32_t *totals_dev, int32_t *pad, int64_t maxk, void *workspace,
                               size_t workspace_bytes, pcs_stream_t stream);
int pcs_search_multi_fill_pad(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                              int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min,
                              const int32_t *lens, const int32_t *begs, const int32_t *pad, int64_t maxk,
                              int32_t *const *idxs, int32_t *const *cens, void *workspace, size_t workspace_bytes,
                              pcs_stream_t stream);

/* ==== libPointUtil replacements: the calls the reference makes either side of the hot path.
 * libPointUtil is an UN-VENDORED dependency of the reference (no source, no pinned version): semantics are restated
 * from the call sites cited below and fixed here; parity is pinned by the numpy oracle only.                      */

/* ---- radius search of a query set into a source set.
 * replaces: libPointUtil.findNeighborRadiusCPU(xyzs, q_idxs, radius) (s3dis_util.py:66: neighbours, in the whole cloud,
 * of the down-sampled points).  Row q = the sources with sq_min < |s - q|^2 < sq_max (strict; sq_min only when
 * use_min), ascending source index.  Two-phase like pcs_search_neighborhood_count/fill: count writes lens[qn],
 * begs[qn] and the edge total (device int32), fill writes idxs[en], cens[en] (cens = query index).              */
size_t pcs_cross_search_workspace_bytes(int64_t sn, int64_t qn);
int pcs_cross_search_count(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, float sq_min, float sq_max,
                           int use_min, int32_t *lens, int32_t *begs, int32_t *total_dev, void *workspace,
                           size_t workspace_bytes, pcs_stream_t stream);
int pcs_cross_search_fill(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, float sq_min, float sq_max,
                          int use_min, const int32_t *begs, int32_t *idxs, int32_t *cens, void *workspace,
                          size_t workspace_bytes, pcs_stream_t stream);

/* ---- exact K nearest sources of every query.
 * replaces: libPointUtil.findNeighborInAnotherCPU(sxyzs, qxyzs, K) (interpolate.py:136, K = 6;
 * interpolate_semantic3d_dense.py:88, K = 8).  idxs[qn, K] ascending by (squared distance, source index), squared
 * distance = fma(dz,dz, fma(dx,dx, dy*dy)) in fp32; dists[qn, K] (nullable) the squared distances.  1 <= K <= 32 <= sn.
 * r0 = first search radius (> 0; a good value holds ~K sources), doubled `rounds` times before the unbounded round;
 * rad2 = caller scratch of qn floats; workspace >= pcs_cross_search_workspace_bytes(sn, 0).  No host sync.        */
int pcs_knn_cross(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, int64_t K, float r0, int64_t rounds,
                  int32_t *idxs, float *dists, float *rad2, void *workspace, size_t workspace_bytes, pcs_stream_t stream);

/* ---- label interpolation.
 * replaces: libPointUtil.interpolateProbsGPU(sxyzs, qxyzs, sprobs, nidxs, nidxs_lens, nidxs_bgs, ratio)
 * (interpolate.py:140, ratio = 1/(2*0.075^2)).  qprobs[q, :] = sum_k w_k sprobs[n_k, :] / sum_k w_k over row q of the
 * CSR, w_k = exp(-|q - s_{n_k}|^2 * ratio); a row whose weights all underflow takes the plain mean; an empty row 0.   */
int pcs_interpolate_probs(const float *sxyzs, const float *qxyzs, const float *sprobs, const int32_t *nidxs,
                          const int32_t *lens, const int32_t *bgs, int64_t qn, int64_t classes, float ratio,
                          float *qprobs, pcs_stream_t stream);

/* ---- covariance features.
 * replaces: libPointUtil.computeCovarsGPU(xyzs, nidxs, nidxs_lens, nidxs_bgs) (s3dis_util.py:72).  covars[q, 9] = the
 * 3x3 covariance of the neighbour coordinates of row q about their mean (row-major), scaled to unit Frobenius norm;
 * zeros when the row has < 2 distinct points.                                                                     */
int pcs_compute_covars(const float *xyzs, const int32_t *nidxs, const int32_t *lens, const int32_t *bgs, int64_t qn,
                       float *covars, pcs_stream_t stream);

/* ---- block cropping.
 * replaces: libPointUtil.gatherBlockPointsGPU(xyz, beg_list, block_size) (aug_util.py:99) = for every block b the
 * points with begs[b] <= (x, y) < begs[b] + block_size, ascending point index (np.nonzero, aug_util.py:75-80).
 * count writes tile_counts[nb, n_tiles] (n_tiles = pcs_block_members_tiles(pn)); the caller forms the exclusive scan
 * of that table in (block, tile) order as int64 tile_offs; fill writes the indices.                                */
int64_t pcs_block_members_tiles(int64_t pn);
int pcs_block_members_count(const float *xyzs, int64_t pn, const float *block_begs, int64_t nb, float block_size,
                            int32_t *tile_counts, pcs_stream_t stream);
int pcs_block_members_fill(const float *xyzs, int64_t pn, const float *block_begs, int64_t nb, float block_size,
                           const int64_t *tile_offs, int32_t *idxs, pcs_stream_t stream);

/* ---- helper: exclusive prefix sum of int32 (thrust::exclusive_scan call sites
 * SearchNeighborhood.cu:143, ComputeRepermutationInfo.cu:75); total_dev nullable.               */
size_t pcs_exclusive_scan_workspace_bytes(int64_t n);
int pcs_exclusive_scan_i32(const int32_t *in, int64_t n, int32_t *out, int32_t *total_dev,
                           void *workspace, size_t workspace_bytes, pcs_stream_t stream);

/* ---- helper: the same exclusive prefix sum for callers that cannot allocate a workspace (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The scan's scratch is a file-scope `static __device__` array (256 KB, held by
 * every GPU that loads the library until exit), so:
 *   - n <= pcs_exclusive_scan_static_max_n() (4 194 304); a larger n returns 1 with a message in pcs_last_error()
 *     and launches nothing — use pcs_exclusive_scan_i32 with a caller workspace there;
 *   - the call performs no cudaMalloc / cudaFree / allocating library call; in, out, total_dev (nullable) are the
 *     caller's device buffers;
 *   - all calls on a device share the one array and it keeps the last call's contents; the scan re-initialises what
 *     it uses, so consecutive calls on one stream are safe, but calls that OVERLAP on the GPU (two streams, two host
 *     threads) race on it and must be serialised by the caller.                                              */
int64_t pcs_exclusive_scan_static_max_n(void);
int pcs_exclusive_scan_i32_static(const int32_t *in, int64_t n, int32_t *out, int32_t *total_dev, pcs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PCSEG_OPS_H */
