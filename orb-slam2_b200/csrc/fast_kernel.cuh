// fast_kernel.cuh -- E3+E4  fast_cells: one thread block per 30-px grid cell of any level
// (src/ORBextractor.cc:953-1006: cell loop, cv::FAST(roi, iniThFAST, nms=true), retry with minThFAST when empty).
//
// cv::FAST semantics used (derived and pinned against cv2 in tests/test_oracle_pin.py):
//   score S(p) = max over the 16 contiguous 9-arcs of min|v - ring| - 1 (threshold independent);
//   corner at th  <=>  S >= th;   NMS keeps p iff S(p) > S(q) for the 8 neighbours q, where q outside the ROI
//   interior or not a corner at th counts 0.  For S(p) >= th the masked comparison equals the raw one, so the
//   survivors at iniTh are the local maxima with S >= iniTh and the minTh retry (taken only when that set is
//   empty) yields the local maxima with S >= minTh -- one scoring pass serves both thresholds.
//   ROI interiors tile [19, W-19) x [19, H-19) without overlap: every pixel is scored exactly once.
//
// sm_100a formulation (integer-ALU / issue bound stage):
//   * two pixels per thread in packed 16-bit lanes: VIMNMX3.S16x2 is a native instruction on sm_100a.
//   * the arc minimum is translation invariant: max_arcs min9(v - ring) = v - min_arcs max9(ring), so the network
//     runs on raw ring values (no 16 subtractions):  A = max_arcs min9(ring), B = min_arcs max9(ring),
//     strength M = max(A - v, v - B); corner at t <=> M > t; score = M - 1.
//     (The negation is applied once, outside the min/max chains: ptxas 12.9 drops a negation folded into
//     VIMNMX3 -- see DESIGN.md.)
//   * the ROI is staged in shared memory widened to 16 bits, twice: copy U holds the pixel pairs that start at even
//     chunk columns, copy S the pairs that start at odd columns (both written with 128-bit stores), so all 17
//     packed reads of a pixel pair are aligned conflict-free LDS.32 with no byte shuffling; which copy serves the
//     even ring offsets depends on the parity of the cell's first interior column.
//   * a warp-uniform quick test on the 4 compass ring pixels skips flat regions.
//   * NMS is separable and stays packed: the scoring pass exchanges scores with the two neighbouring lanes
//     (2 shuffles), keeps s' = s where s beats its horizontal neighbours and the 3-wide row maximum m3, and stores
//     one word {s'0, s'1, m3_0, m3_1} per pixel pair; the second pass needs the words above and below only.
//   * each warp owns a contiguous band of rows and appends its survivors to its own shared-memory list in lane
//     order (= the reference's row-major order); after the block-wide iniTh/minTh vote the lists are copied to the
//     cell's fixed slot -- no atomics, no block scan.
#pragma once
#include "extract_kernels.cuh"

namespace orbb200 {

#ifdef ORBB200_EXTRACT_IMPL

constexpr int FC_THREADS = 128, FC_WARPS = 4;
constexpr int FC_SPW = 48;      // staged row pitch in 32-bit words per copy (96 pixels = 6 chunks of 16)
constexpr int FC_MAXROWS = 66;  // hCell + 6 <= 66 (hCell < 60 because nRows = floor(height / 30))
constexpr int FC_LISTCAP = 256; // survivors per warp: <= ceil(15 / 2) * 30 strict local maxima in its <= 15 rows

__host__ __device__ inline size_t fast_cells_smem(int rows)
{
    // two widened copies + one score word per pixel pair (rows - 6 interior rows rounded up to a group, 2 border rows,
    // 32 words per row at most) + the per-warp survivor lists
    return (size_t)rows * FC_SPW * 4 * 2 + (size_t)(rows + 2) * 32 * 4 + (size_t)FC_WARPS * FC_LISTCAP * 4;
}

__device__ __forceinline__ unsigned int widen_lo(unsigned int w) { return __byte_perm(w, 0u, 0x4140); }
__device__ __forceinline__ unsigned int widen_hi(unsigned int w) { return __byte_perm(w, 0u, 0x4342); }
// a - b per 16-bit half for halves in [0, 0x7FFF], BIASED by 0x8000: one plain 32-bit subtraction, no borrow can cross
// the halves, bit 15 of a half is clear <=> a < b there; biased values order like unsigned halves (VIMNMX.U16x2).
// (Deliberately not __vsub2: nvcc 12.9 splits its 16x2 negation constant wrongly in a peeled loop iteration --
// the low half lost its +1 -- see DESIGN.md.)
__device__ __forceinline__ unsigned int bsub2(unsigned int a, unsigned int b) { return (a | 0x80008000u) - b; }
// 0xFFFF in each 16-bit half whose bit 15 is set (prmt replicates the msb of bytes 1 and 3)
__device__ __forceinline__ unsigned int neg_mask2(unsigned int w)
{
    unsigned int r;
    asm("prmt.b32 %0, %1, 0, 0xBB99;" : "=r"(r) : "r"(w));
    return r;
}


#endif // ORBB200_EXTRACT_IMPL

} // namespace orbb200
