// ref_shim.cu — TEST INFRASTRUCTURE.  extern "C" doorway onto the REFERENCE's own host launchers.
// The reference .cu files are #included from where they lie (-I$(REF)/tf_ops, see Makefile); no
// reference source is copied into this repo.  Output: oracle/_ref/libref.so (git-ignored), used on
// the GPU box by tests/ (parity pinning) and by bench.py (the "reference CUDA on the same B200"
// column).  Never loaded by the product package.
//
// One addition, SURVEY note N1: the reference declares, instantiates and calls
// neighborSumFeatGatherGPU (TFNeighborKernel.h:84-93, TFNeighborKernelNew.cu:394,
// TFNeighborSumFeatGather.cc:66) but ships no definition of it.  It is supplied below exactly in the
// style of its siblings (memset because the kernel accumulates with +=, then the shared 2-D launch
// heuristic of TFNeighborKernelNew.cu:152-171).
#include "TFNeighborKernel.h"
#include "TFCudaCommon.h"

template <typename FLT_TYPE, typename INT_TYPE>
__global__ void sumFeatGather(FLT_TYPE *ifeats, INT_TYPE *nidxs_lens, INT_TYPE *nidxs_bgs, INT_TYPE fd,
                              INT_TYPE pn, FLT_TYPE *ogfeats_sum);

template <typename FLT_TYPE, typename INT_TYPE>
void neighborSumFeatGatherGPU(FLT_TYPE *d_ifeats, INT_TYPE *d_nidxs_lens, INT_TYPE *d_nidxs_bgs, INT_TYPE pn,
                              INT_TYPE fd, INT_TYPE csum, FLT_TYPE *d_ogfeats_sum) {
    int tdim0, tdim1, tdim2 = 1;
    int bdim0, bdim1, bdim2 = 1;
    tdim1 = 1024 / (tdim2);
    if (fd < tdim1) tdim1 = infTwoExp(fd);
    bdim1 = fd / tdim1;
    if (fd % tdim1 > 0) bdim1++;
    tdim0 = 1024 / (tdim1 * tdim2);
    if (pn < tdim0) tdim0 = infTwoExp(pn);
    bdim0 = pn / tdim0;
    if (pn % tdim0 > 0) bdim0++;
    dim3 block_dim(bdim0, bdim1, bdim2);
    dim3 thread_dim(tdim0, tdim1, tdim2);
    gpuErrchk(cudaMemset(d_ogfeats_sum, 0, pn * fd * sizeof(FLT_TYPE)))
    sumFeatGather<FLT_TYPE, INT_TYPE><<<block_dim, thread_dim>>>(d_ifeats, d_nidxs_lens, d_nidxs_bgs, fd, pn,
                                                                  d_ogfeats_sum);
    gpuErrchk(cudaGetLastError())
}

#include "TFNeighborKernel.cu"
#include "TFNeighborKernelNew.cu"
#include "PoolingIndex/ComputeVoxelIdx.cu"
#include "PoolingIndex/ComputePermutationInfo.cu"
#include "PoolingIndex/ComputeRepermutationInfo.cu"
#include "PoolingIndex/PermutateFeature.cu"
#include "PoolingIndex/ComputeDiffXYZ.cu"
#include "PoolingIndex/SearchNeighborhood.cu"
#include "PoolingIndex/SearchNeighborhoodFixed.cu"

typedef unsigned int u32;
extern "C" {
void ref_sync() { cudaDeviceSynchronize(); }
void ref_neighbor_scatter(float *ifeats, u32 *nidxs, u32 *lens, u32 *bgs, u32 pn, u32 ifn, float *sfeats, int use_diff) {
    neighborScatterGPU<float, u32>(ifeats, nidxs, lens, bgs, pn, ifn, sfeats, use_diff != 0);
}
void ref_neighbor_gather(float *sfeats, u32 *nidxs, u32 *lens, u32 *bgs, u32 pn, u32 ifn, float *ifeats, int use_diff) {
    neighborGatherGPU<float, u32>(sfeats, nidxs, lens, bgs, pn, ifn, ifeats, use_diff != 0);
}
void ref_sum_feat_gather(float *ifeats, u32 *lens, u32 *bgs, u32 pn, u32 fd, u32 csum, float *out) {
    neighborSumFeatGatherGPU<float, u32>(ifeats, lens, bgs, pn, fd, csum, out);
}
void ref_sum_feat_scatter(float *in, u32 *cidxs, u32 pn, u32 fd, u32 csum, float *out) {
    neighborSumFeatScatterGPU<float, u32>(in, cidxs, pn, fd, csum, out);
}
void ref_max_feat_gather(float *ifeats, u32 *vlens, u32 *vbegs, u32 pn2, u32 fd, float *g, u32 *max_idxs) {
    neighborMaxFeatGatherGPU<float, u32>(ifeats, vlens, vbegs, pn2, fd, g, max_idxs);
}
void ref_max_feat_scatter(float *g, u32 *max_idxs, u32 *vbegs, u32 *vlens, u32 pn1, u32 pn2, u32 fd, float *out) {
    neighborMaxFeatScatterGPU<float, u32>(g, max_idxs, vbegs, vlens, pn1, pn2, fd, out);
}
void ref_locw_sum(float *lw, u32 *lens, u32 *bgs, u32 pn, u32 m, float *out) {
    locWSumForwardGPU<float, u32>(lw, lens, bgs, pn, m, out);
}
void ref_locw_sum_backward(float *dsum, u32 *lens, u32 *bgs, u32 csum, u32 pn, u32 m, float *dlw) {
    locWSumBackwardGPU<float, u32>(dsum, lens, bgs, csum, pn, m, dlw);
}
void ref_locw_feat_sum(float *tfeats, float *lw, u32 *lens, u32 *bgs, u32 pn, u32 ofn, u32 m, float *out) {
    locWFeatSumForwardGPU<float, u32>(tfeats, lw, lens, bgs, pn, ofn, m, out);
}
void ref_locw_feat_sum_backward(float *tfeats, float *lw, float *dsum, u32 *lens, u32 *bgs, u32 csum, u32 pn,
                                u32 ofn, u32 m, float *dtfeats, float *dlw) {
    locWFeatSumBackwardGPU<float, u32>(tfeats, lw, dsum, lens, bgs, csum, pn, ofn, m, dtfeats, dlw);
}
void ref_eliminate_center(u32 *in, u32 *lens, u32 *bgs, u32 pn, u32 *on, u32 *ol, u32 *ob, u32 *oc) {
    eliminateCenterGPU<u32>(in, lens, bgs, pn, on, ol, ob, oc);
}
void ref_voxel_index(float *pts, u32 *out, int pn, int stride, float len, float mx, float my, float mz) {
    computeVoxelIdxImpl(pts, out, pn, stride, len, mx, my, mz);
}
// returns vn; lens/begs/cens are copied into caller buffers (sized pn) and the thrust-side
// allocations freed, like ComputePermutationInfo.cc:157-159 does through copyDataAndFree.
int ref_permutation_info(int *vidx, int *classes, int *o2p, int *p2o, int *lens, int *begs, int *cens, int pn) {
    int *l, *b, *c, vn;
    if (classes) computePermutationInfoWithClassesImpl(vidx, classes, o2p, p2o, &l, &b, &c, &vn, pn);
    else computePermutationInfoImpl(vidx, o2p, p2o, &l, &b, &c, &vn, pn);
    copyDataAndFree<int>(l, lens, vn);
    copyDataAndFree<int>(b, begs, vn);
    copyDataAndFree<int>(c, cens, pn);
    return vn;
}
void ref_repermutation_info(int *o2p2, int *lens, int *begs, int *rlens, int *rbegs, int *rcens, int *ro2p1,
                            int *rp2o1, int pn1, int pn2) {
    computeRepermutationInfoImpl(o2p2, lens, begs, rlens, rbegs, rcens, ro2p1, rp2o1, pn1, pn2);
}
void ref_permutate_feature(float *feats, int *idxs, float *out, int pn, int ps) {
    permutateFeature<float>(feats, idxs, out, pn, ps);
}
void ref_diff_xyz(float *xyzs, float *cxyzs, float *dxyzs, int *cidxs, int pn1, int pn2) {
    computeDiffXYZImpl(xyzs, cxyzs, dxyzs, cidxs, pn1, pn2);
}
int ref_search_count(float *xyzs, int *lens, int *begs, float sq_min, float sq_max, int use_min, int pn) {
    return use_min ? searchNeighborhoodCountRangeImpl(xyzs, lens, begs, sq_min, sq_max, pn)
                   : searchNeighborhoodCountImpl(xyzs, lens, begs, sq_max, pn);
}
void ref_search_fill(float *xyzs, int *idxs, int *cens, int *begs, float sq_min, float sq_max, int use_min, int pn) {
    if (use_min) searchNeighborhoodRangeImpl(xyzs, idxs, cens, begs, sq_min, sq_max, pn);
    else searchNeighborhoodImpl(xyzs, idxs, cens, begs, sq_max, pn);
}
void ref_search_fixed(float *xyzs, int *idxs, int *cens, int *lens, int *begs, float sq_min, float sq_max,
                      int use_min, int K, int pn) {
    if (use_min) searchNeighborhoodFixedRangeImpl(xyzs, idxs, cens, sq_min, sq_max, K, pn);
    else searchNeighborhoodFixedImpl(xyzs, idxs, cens, sq_max, K, pn);
    computeNeighborhoodStats(lens, begs, K, pn);
}
}
