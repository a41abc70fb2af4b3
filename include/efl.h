/* efl.h -- C ABI of libefl.so: B200 (sm_100a) kernels for the EF_layers hot path of Isomorpfishm/SINGA.
 *
 * The reference has no FFI layer: its boundary for this path is the Python surface of model/EF_layers.py as
 * consumed by model/Embedding.py:12-33.  singa_b200/EF_layers.py keeps that surface and calls the functions
 * below (through ctypes, singa_b200/_lib.py).  Each entry cites the reference code it replaces
 * (paths relative to the reference root, model/EF_layers.py unless stated).
 *
 * Conventions: C linkage, POD arguments only; every pointer is a DEVICE pointer borrowed for the call (the
 * row-map arrays of the S2 entry points are HOST arrays); no hidden allocation -- workspaces are supplied by the
 * caller; all work is enqueued on `stream` (a cudaStream_t passed as void*), no synchronisation inside;
 * return 0 on success, otherwise a cudaError_t value or an EFL_ERR_* code, with a message available from
 * efl_last_error_string() (thread-local).  Shapes use L=(lmax+1)^2 coefficients, M kept coefficients with
 * |m|<=mmax, WSZ = sum_l min(2l+1, 2mmax+1)*pad4(2l+1) Wigner floats per edge (rows padded to 16 bytes); built for (lmax,mmax) in
 * {(6,2), (8,2), (4,2)}.
 *
 * Edge operand layout ("m-primary, split by order"): for order m a dense [rows_m*Ec, (lmax+1-m)*CH] matrix,
 * rows_m = 1 for m=0, else 2 (row 2e = +m "real", row 2e+1 = -m "imag"), Ec = edges of the chunk, in
 * destination-sorted order.  The l<->m permutation einsums of the reference (:354-359) are folded into these
 * layouts.  "dup layout" for gradients w.r.t. SO2_m_Convolution outputs: [2Ec, 2*half] with interleaved
 * columns, row 2e = (d re_q, d im_q), row 2e+1 = (d im_q, -d re_q), matching weights whose rows are
 * interleaved (real_q, imag_q).
 */
#ifndef EFL_H_
#define EFL_H_

#ifdef __cplusplus
extern "C" {
#endif

#define EFL_ERR_UNSUPPORTED 10001
#define EFL_ERR_ARG 10002

int efl_version(void);
const char* efl_last_error_string(void);
int efl_device_cc(int device, int* major, int* minor);

/* SO3_Rotation.set_wigner / RotationToWignerDMatrix / wigner_D / _z_rot_mat (:485-528, :2207-2229).
 * rot [E,3,3] fp32 edge frames, jd = concatenated J_l (l=0..lmax, row-major; the Jd.pt matrices, a device pointer:
 * copied stream-ordered into the kernel's constant bank on every call) -> wig [E, WSZ]: the rows |m'| <= mmax of every
 * D^l, each padded to 16 bytes (WSZ = 296 floats for (6,2), 476 for (8,2)).  Built for mmax = 2, lmax in {4, 6, 8}. */
int efl_wigner_from_rotmat(int lmax, int mmax, const float* rot, const float* jd, float* wig, int E, void* stream);

/* Edge operand layout shared by the entry points below ("edge rows"): every edge of a chunk owns ONE row of
 * `pitch` elements (fp32 or bf16); inside the row the kept coefficients are in m-primary order, the m = 0 block
 * ((lmax+1) degrees x CH channels), then for m = 1, 2 the +m block followed by the -m block ((lmax+1-m) x CH each).
 * The three pointers a0/a1/a2 (v0.., dy0..) address the start of the m = 0 / 1 / 2 blocks of the chunk's first edge.
 * (+m | -m) adjacent makes an edge row the [x_re | x_im] operand of SO2_m_Convolution written as ONE real GEMM
 * (:721-729: out_re = x_re Wr^T - x_im Wi^T, out_im = x_im Wr^T + x_re Wi^T  ==  [x_re | x_im] [[Wr,-Wi],[Wi,Wr]]^T),
 * so forward, dgrad and wgrad of all six SO(2) linears are plain GEMMs over M = edges with no duplicated layout. */

/* Gather + rotate-in: SO3_Embedding.clone/_expand_edge + cat (:1110-1119), SO3_Rotation.rotate (:494-497),
 * _m_primary (:354-355) and the radial multiply of SO2_Convolution (:820-850).
 * x_src [Ns,L,C], x_dst [Nt,L,C]; src/dst/perm int32 [E] (dst-sorted order; perm = original edge id);
 * wig [E,WSZ] and rad [E,rad_ld] in ORIGINAL edge order (rad may be NULL); a0..a2 + pitch: edge rows (CH = 2C)
 * for chunk edges [e0, e0+Ec).  ones_pad > 0: the ones_pad columns in front of the m = 0 block are set to
 * (1, 0, ..., 0): the bias column of the so2_conv_1 m = 0 GEMM (fc_m0.bias, :820-825). */
int efl_rotate_in_fwd(int lmax, int mmax, int C, int bf16_out, const float* x_src, const float* x_dst,
                      const int* src, const int* dst, const int* perm, const float* wig, const float* rad,
                      int rad_ld, void* a0, void* a1, void* a2, long long pitch, int ones_pad, int e0, int Ec,
                      void* stream);
/* Backward of the above: da0..da2 (dgrad of so2_conv_1, edge rows) -> drad [E,rad_ld] (original order) and per-edge
 * d x [E, L, 2C] (sorted order; columns [0,C) source half, [C,2C) destination half). */
int efl_rotate_in_bwd(int lmax, int mmax, int C, int bf16_in, const float* x_src, const float* x_dst,
                      const int* src, const int* dst, const int* perm, const float* wig, const float* rad,
                      int rad_ld, const void* da0, const void* da1, const void* da2, long long pitch, float* drad,
                      float* dxe, int e0, int Ec, void* stream);
/* Node-owned variant: one warp per node sums d x over its edges in registers (fixed order) and writes
 * dx_out [N,L,C] once; the per-edge d x tensor never exists.  side 0 = destination half: seg_ptr = CSR rowptr,
 * seg_idx NULL, nodes [n0,n1) of the chunk.  side 1 = source half: seg_ptr/seg_idx = source grouping (sorted
 * positions, ascending per source), all source nodes, accumulate = 1 from the second chunk on.  drad as above. */
int efl_rotate_in_bwd_nodes(int lmax, int mmax, int C, int bf16_in, int side, const float* x_node,
                            const int* seg_ptr, const int* seg_idx, const int* perm, const float* wig,
                            const float* rad, int rad_ld, const void* da0, const void* da1, const void* da2,
                            long long pitch, float* drad, float* dx_out, int e0, int Ec, int n0, int n1,
                            int accumulate, void* stream);
/* Deterministic backward of the gather (:326-328): out[n,i,c] = sum_{k in [ptr[n],ptr[n+1])}
 * buf[idx?idx[k]:k, i, coff+c]; buf rows are L*ldc floats. */
int efl_segment_sum_rows(const float* buf, int L, int ldc, int coff, int C, const int* ptr, const int* idx,
                         float* out, int num_nodes, int accumulate, void* stream);

/* alpha*value (:1186-1190) + SO3_Rotation.rotate_inv with get_rotate_inv_rescale (:500-505, :1530-1549) +
 * SO3_Embedding._reduce_edge (:342-351) fused: out [Nt, L, heads*value] for destinations [n0,n1), one CTA per
 * destination, fixed summation order, no atomics.  v0..v2 + pitch: value edge rows (CH = heads*value);
 * alpha [E,heads] sorted order (NULL = 1). */
int efl_rotate_back_aggregate_fwd(int lmax, int mmax, int heads, int value_ch, int bf16_in, const void* v0,
                                  const void* v1, const void* v2, long long pitch, const float* alpha,
                                  const int* rowptr, const int* perm, const float* wig, float* out, int n0, int n1,
                                  int e0, void* stream);
/* Backward: dout [Nt,L,heads*value] -> dy0..dy2 + dy_pitch (d value, edge rows like v) and dalpha
 * [efl_rotate_back_parts(lmax,mmax), E, heads]: one partial plane per degree group (planes part_stride floats
 * apart), summed in fixed order by efl_segment_softmax_bwd. */
int efl_rotate_back_parts(int lmax, int mmax);
int efl_rotate_back_aggregate_bwd(int lmax, int mmax, int heads, int value_ch, int bf16, const void* v0,
                                  const void* v1, const void* v2, long long pitch, void* dy0, void* dy1, void* dy2,
                                  long long dy_pitch, const float* alpha, const int* rowptr, const int* perm,
                                  const float* wig, const float* dout, float* dalpha, long long part_stride, int n0,
                                  int n1, int e0, void* stream);

/* EdgeDegreeEmbedding tail (:116-147): m0 [E,(lmax+1)*C] (rad_func output, original order) -> out [Nt,L,C]. */
int efl_edge_degree_fwd(int lmax, int mmax, int C, const float* m0, const int* rowptr, const int* perm,
                        const float* wig, float rescale_factor, float* out, int Nt, void* stream);
int efl_edge_degree_bwd(int lmax, int mmax, int C, const float* dout, const int* dst_orig, const float* wig,
                        float rescale_factor, float* dm0, int E, void* stream);

/* Attention logits (:1175-1178): extra[:, :heads*32] -> LayerNorm(32) (:1024) -> SmoothLeakyReLU (:1669-1677)
 * -> dot alpha_dot -> logits [Ec, heads]. */
int efl_alpha_logits_fwd(const float* extra, int ld, int heads, int alpha_ch, const float* ln_w, const float* ln_b,
                         const float* alpha_dot, float* logits, int Ec, void* stream);
/* Backward; part [nblocks, heads*32+64] receives per-block partial sums of (d alpha_dot | d ln_w | d ln_b). */
int efl_alpha_logits_bwd(const float* extra, int ld, int heads, int alpha_ch, const float* ln_w, const float* ln_b,
                         const float* alpha_dot, const float* dlogits, void* dextra, int ldo, int bf16_out,
                         float* part, int nblocks, int Ec, void* stream);

/* torch_geometric.utils.softmax(alpha, edge_index[1]) (:1180; torch_geometric 2.3.1) over the dst-sorted CSR:
 * exp(x - max_seg) / (sum_seg + 1e-16), one warp per destination in [n0,n1). */
int efl_segment_softmax_fwd(const float* logits, int heads, const int* rowptr, float* alpha, int n0, int n1,
                            void* stream);
/* dalpha: nparts planes [E,heads], part_stride floats apart (nparts = 1 for a plain gradient). */
int efl_segment_softmax_bwd(const float* alpha, const float* dalpha, int nparts, long long part_stride, int heads,
                            const int* rowptr, float* dlogits, int n0, int n1, void* stream);

/* SeparableS2Activation / S2Activation (:1736-1773) with the SO3_Grid matrices (:531-621) given as
 * to_grid, from_grid [G, M] (G = res_beta*res_alpha).  Row maps (HOST arrays): nseg segments, base[nseg]
 * device pointers, stride[nseg] elements per item, first[nseg+1] row ranges.  gate [n_items, gate_ld]. */
int efl_s2act_fwd(int M, int G, int CH, int bf16_in, int bf16_out, int in_nseg, void* const* in_base,
                  const long long* in_stride, const int* in_first, int out_nseg, void* const* out_base,
                  const long long* out_stride, const int* out_first, const float* gate, long long gate_ld,
                  const float* to_grid, const float* from_grid, long long n_items, void* stream);
int efl_s2act_bwd(int M, int G, int CH, int bf16_in, int bf16_d, int in_nseg, void* const* in_base,
                  const long long* in_stride, const int* in_first, int din_nseg, void* const* din_base,
                  const long long* din_stride, const int* din_first, int dout_nseg, void* const* dout_base,
                  const long long* dout_stride, const int* dout_first, int dup, const float* gate,
                  long long gate_ld, void* dgate, long long dgate_ld, const float* to_grid, const float* from_grid,
                  long long n_items, void* stream);

/* SO2_Convolution / SO2_m_Convolution linear maps (:807-875, :721-729), fp32 SIMT accuracy mode.
 * NT: C[i,n] = sum_k A[i,k] B[n,k] (+bias).  epilogue 0 plain; 1 split (cols < nsplit -> C0 fp32, rest -> C);
 * 2 complex combine on row pairs / interleaved column pairs -> C [M, N/2]. */
int efl_gemm_nt_f32(int epilogue, const float* A, long long lda, const float* B, long long ldb, const float* bias,
                    float* C, long long ldc, float* C0, long long ldc0, int nsplit, int M, int N, int K,
                    void* stream);
/* TN (weight gradients): C[n,k] (+)= sum_i A[i,n] B[i,k], deterministic split over i.
 * workspace: efl_gemm_tn_splits(Mred) * N1 * N2 floats. */
int efl_gemm_tn_splits(int Mred);
int efl_gemm_tn_f32(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc,
                    int accumulate, float* workspace, int Mred, int N1, int N2, void* stream);
/* Bias gradients: out[c] (+)= sum_r X[r,c]; workspace (rows/512 + 2) * cols floats. */
int efl_colsum(int bf16, const void* X, long long ld, int rows, int cols, float* out, int accumulate,
               float* workspace, void* stream);
/* efl_colsum without the workspace argument, for callers that are not able to allocate device memory: a process whose
 * memory pool is sized once and then closed, a plugin working inside buffers its host program gave it, or code inside
 * a stream capture, where cudaMalloc fails.  No counterpart in the reference (there every op allocates freely through
 * torch); it exists only because of how this library is embedded.  The partial sums go to an array that is part of
 * libefl.so itself: EFL_COLSUM_STATIC_MAX_FLOATS floats (4 MiB) held on every GPU that loads the library, until the
 * process exits.  Shapes with (ceil(rows/512) + 1) * cols above that limit return EFL_ERR_ARG (message in
 * efl_last_error_string()); nothing is allocated or freed by this call.  Results are bit-identical to efl_colsum.
 * The first call on a device resolves the array's address on that device: make it with that device current and
 * before any stream capture starts.  `stream` must be a stream of the current device.  All calls on a GPU share the
 * array, which keeps the previous call's contents (every slot that is read is rewritten first, so no clearing is
 * needed); calls that can overlap on the GPU race on it -- keep them on one stream or order them with events. */
#define EFL_COLSUM_STATIC_MAX_FLOATS (1 << 20)
int efl_colsum_static(int bf16, const void* X, long long ld, int rows, int cols, float* out, int accumulate,
                      void* stream);

/* The same linear maps on the 5th-generation tensor cores: bf16 operands, fp32 accumulation in TMEM
 * (tcgen05.mma), operands staged by TMA, persistent warp-specialised CTAs.  Same epilogues as
 * efl_gemm_nt_f32; C is bf16 (C0 fp32).  N (and N2) must be multiples of 16; row pitches multiples of 8. */
int efl_tc_gemm_nt(int epilogue, const void* A, long long lda, const void* B, long long ldb, const float* bias,
                   void* C, long long ldc, float* C0, long long ldc0, int nsplit, int M, int N, int K, void* stream);
int efl_tc_gemm_tn_splits(int Mred, int N1, int N2);
/* CTA-pair (cta_group::2, 256-row tiles) variant of efl_tc_gemm_nt: 0 off (default), 1 large problems, 2 always.
 * Returns the previous mode; a mode outside 0..2 only queries. */
int efl_tc_gemm_pair_mode(int mode);
/* Operand ring size of the tcgen05 GEMM CTAs in KB (96..192, default 192 or EFL_GEMM_RING_KB): a smaller ring leaves
 * shared memory for HBM-bound kernels of another stream to co-reside with the persistent GEMM.  Returns the previous value. */
int efl_tc_gemm_ring_kb(int kb);
int efl_tc_gemm_tn(const void* A, long long lda, const void* B, long long ldb, float* C, long long ldc,
                   int accumulate, float* workspace, int Mred, int N1, int N2, void* stream);

/* Operand kernels of the node-level S2 activation on grids outside the fused tcgen05 node kernel (lmax = 8, config 5;
 * FeedForwardNetwork :234-270 with SeparableS2Activation :1757-1773 as two plain GEMMs over (node, channel) rows):
 * pack x [N,L,CH] fp32 -> bf16 rows [(n,c), LP] (zero padded, optionally l = 0 zeroed); unpack back to fp32 with the
 * l = 0 row replaced by SiLU(gate [N,CH]) when gate != NULL; efl_bf16_silu: a <- SiLU(a) (mode 0) or
 * a <- a * SiLU'(b) (mode 1) in place on bf16 arrays. */
int efl_node_rows_pack(const float* x, int N, int L, int CH, int LP, int zero_row0, void* out, void* stream);
int efl_node_rows_unpack(const void* yt, int N, int L, int CH, int LP, const float* gate, float* y, void* stream);
int efl_bf16_silu(void* a, const void* b, long long n, int mode, void* stream);

/* bf16x3 operand preparation (accurate tensor-core mode): fp32 x [rows, cols] -> three bf16 copies (hi, hi, lo) for an
 * A operand or (hi, lo, hi) for a B operand (second = 1), side by side along the columns (stack = 0, NT GEMMs) or one
 * under the other (stack = 1, TN GEMMs), so that ONE efl_tc_gemm_* launch with a 3x longer reduction computes
 * a_hi b_hi + a_hi b_lo + a_lo b_hi in fp32 (csrc/split3.cu). */
int efl_split3(const float* x, long long ld, long long rows, int cols, void* out, long long ldo, int stack, int second,
               void* stream);

/* SeparableS2Activation of the edge path (:1757-1773) on the tensor cores, bf16 buffers: per edge two (fwd) or
 * three (bwd) tcgen05 MMAs with the channel axis as M; the grid values live only in TMEM.  h: hidden edge rows
 * [Ec, M, 128] (dense, m-primary); fwd: out + ld_out = output edge rows; bwd: d = d(out) rows (dense like h), out =
 * first hidden column of the so2_conv_1 dY row and dgate = its gate column, both with row pitch ld_out.
 * T, F: to/from grid matrices [G, M] fp32 with m-primary columns. */
int efl_s2act_tc(int bwd, int lmax, int mmax, int G, const void* h, const void* d, void* out, long long ld_out,
                 void* dgate, const float* gate, long long gate_ld, const float* T, const float* F, int Ec,
                 void* stream);

/* Node-level variant for FeedForwardNetwork (:234-270): x [N, M, cht] bf16 (cht a multiple of 128), gate
 * [N, gate_ld] fp32, T/F [G, M] fp32; fwd: out [N, M, cht]; bwd: dy -> out = d x, dgate [N, cht] (bf16). */
int efl_s2act_tc_nodes(int bwd, int M, int G, int cht, const void* x, const void* dy, void* out, void* dgate,
                       const float* gate, long long gate_ld, const float* T, const float* F, int N, void* stream);

/* SO3_LinearV2 (:624-674), fp32: out[n,i,:] = x[n,i,:] W[l(i)]^T (+bias on i=0); x [N,L,Cin], W [lmax+1,Cout,Cin].
 * bwd_data takes the per-degree transposed weights Wt [lmax+1,Cin,Cout]; bwd_weight needs L*Cout*Cin floats. */
int efl_so3_linear_fwd(const float* x, const float* W, const float* bias, float* out, int N, int lmax, int Cin,
                       int Cout, void* stream);
int efl_so3_linear_bwd_data(const float* dout, const float* Wt, float* dx, int N, int lmax, int Cin, int Cout,
                            void* stream);
int efl_so3_linear_bwd_weight(const float* dout, const float* x, float* dW, float* workspace, int N, int lmax,
                              int Cin, int Cout, void* stream);

/* Operand layout kernels that put SO3_LinearV2 (:624-674) on efl_tc_gemm_nt / efl_tc_gemm_tn.  rows = N*(lmax+1)^2.
 * expand: x [rows, C] fp32 -> xe [rows, KP] bf16; expand=1 puts x[r,:] in column block l(i) (zeros elsewhere) and a 1
 * in bias_col on i=0 rows (bias_col < 0: none); expand=0 is a plain bf16 cast with zero padding.
 * select: y[r, c] = ye[r, l(i)*C + c] (+ bias[c] on i=0), ye fp32 with row pitch ld. */
int efl_so3_expand(const float* x, void* xe, long long rows, int lmax, int C, int KP, int expand, int bias_col,
                   void* stream);
int efl_so3_select(const float* ye, long long ld, const float* bias, float* y, long long rows, int lmax, int C,
                   void* stream);

/* RadialFunction (:1634-1657) for edge_channels_list [K0, 16, 16] + output width NO, fused:
 * Linear(K0,16) -> LayerNorm(16) -> SiLU -> Linear(16,16) -> LayerNorm(16) -> SiLU -> Linear(16,NO).  x [E,K0] -> out [E,NO]
 * fp32.  Weights in the nn.Linear / nn.LayerNorm layouts of RadialFunction.net.{0,1,3,4,6}.  Backward recomputes the
 * hidden activations and writes per-CTA parameter-gradient partials part [nblocks, efl_radial_mlp_part_size(K0,NO)]:
 *   dW3 [NO,16] | db3 [NO] | dW2 [16,16] | db2 | dg2 | dbeta2 | dW1 [16,K0] | db1 | dg1 | dbeta1   (summed by the caller
 * in fixed order: deterministic, no atomics).  K0 in {16,32,48,64}, NO % 4 == 0, NO <= 768. */
int efl_radial_mlp_part_size(int K0, int NO);
int efl_radial_mlp_fwd(const float* x, int K0, const float* W1, const float* b1, const float* g1, const float* be1,
                       const float* W2, const float* b2, const float* g2, const float* be2, const float* W3,
                       const float* b3, int NO, float* out, int E, void* stream);
int efl_radial_mlp_bwd(const float* x, int K0, const float* W1, const float* b1, const float* g1, const float* be1,
                       const float* W2, const float* b2, const float* g2, const float* be2, const float* W3,
                       const float* b3, int NO, const float* dout, float* dx, float* part, int nblocks, int E,
                       void* stream);

/* radius_graph (torch-cluster 1.6.3 radius_graph, un-vendored; reference gen.py:17) as a cell-list
 * neighbour search emitting a destination-sorted CSR: edge (j -> i) iff j != i (unless loop), same batch item and
 * ((xi-xj)^2 + (yi-yj)^2) + (zi-zj)^2 < r*r in fp32 (separately rounded); at most cap sources per target (smallest
 * indices kept); targets ascending, sources ascending inside a target.  pos [N,3] fp32, batch [N] int32 or NULL.
 *  bbox : box_out [B,6] floats (min xyz | max xyz) for the host to size the grid; box_ws B*6 uints scratch.
 *  count: cells of edge cs >= r (nx*ny*nz per batch item, origin [B,3] device floats); writes cell_of [N],
 *         cell_start [ncells+1], sorted [N] float4, deg [2*(N+1)] (full | capped), rowptr_full/rowptr_cap [N+1];
 *         scratch 2*(ncells+1) ints, temp efl_radius_graph_temp_bytes(max(N, ncells) + 1) bytes.
 *  fill : col [rowptr_full[N]] sources, every row sorted ascending; max_deg = largest full degree.
 *  emit : edge_index [2,E] int64, E = rowptr_cap[N]; flow 0: row 0 = sources (source_to_target), 1: swapped. */
int efl_radius_graph_bbox(const float* pos, const int* batch, int N, int nbatch, void* box_ws, float* box_out,
                          void* stream);
long long efl_radius_graph_temp_bytes(long long max_items);
int efl_radius_graph_count(const float* pos, const int* batch, int N, int nbatch, const float* origin, float cs, int nx,
                           int ny, int nz, float r, int loop, int cap, int* cell_of, int* cell_start, int* scratch,
                           void* sorted, int* deg, int* rowptr_full, int* rowptr_cap, void* temp, long long temp_bytes,
                           void* stream);
int efl_radius_graph_fill(int N, int nbatch, float cs, int nx, int ny, int nz, float r, int loop, const int* cell_of,
                          const int* cell_start, const void* sorted, const int* rowptr_full, int* col, int max_deg,
                          void* stream);
int efl_radius_graph_emit(const int* rowptr_full, const int* rowptr_cap, const int* col, int N, long long E,
                          long long* edge_index, int flow_target_to_source, void* stream);

/* knn_graph(pos, k, batch, flow='target_to_source') of torch-cluster 1.6.3 (un-vendored; reference call sites
 * model/CProMG.py:293 and :330) on the same cell list: after efl_radius_graph_bbox, this bins/scans/scatters the points
 * and finds, one warp per point, the k (<= 64) nearest OTHER points of the same batch item (fp32 ((dx^2+dy^2)+dz^2),
 * ties to the smaller index), searching shells of cells until the k-th distance lies inside the covered radius.
 * nbr [N*k] (ascending distance, -1 padded), cnt [N+1], rowptr_full [N+1] (= i*k), rowptr [N+1] = scan(cnt);
 * efl_radius_graph_emit(rowptr_full, rowptr, nbr, N, rowptr[N], edge_index, flow) writes the int64 edge list. */
int efl_knn_graph(const float* pos, const int* batch, int N, int nbatch, const float* origin, float cs, int nx, int ny,
                  int nz, int K, int loop, int* cell_of, int* cell_start, int* scratch, void* sorted, int* nbr,
                  int* cnt, int* rowptr_full, int* rowptr, void* temp, long long temp_bytes, void* stream);

/* Destination-sorted CSR of a caller-supplied edge_index plus the source-side grouping (csrc/csr.cu).  The reference
 * works on the raw edge list (gather EF_layers.py:326-328, index_add_ :342-351, softmax index :1180); every kernel of
 * this library consumes edges grouped by destination instead.  src64/dst64: rows 0 and 1 of edge_index ([E] int64).
 * Outputs (int32): perm [E] sorted position -> original edge id (stable), rowptr [num_dst+1], src_sorted/dst_sorted
 * [E], dst_orig [E] (original order), src_ptr [num_src+1], src_idx [E] sorted positions grouped by source (stable).
 * scratch 3*E ints; temp efl_csr_temp_bytes(E, max(num_src, num_dst)) bytes; *status (device int, zeroed by the
 * caller) becomes 1 when an id is outside [0, num_src) / [0, num_dst). */
long long efl_csr_temp_bytes(long long E, long long max_nodes);
int efl_csr_from_edge_index(const long long* src64, const long long* dst64, int E, int num_src, int num_dst, int* perm,
                            int* rowptr, int* src_sorted, int* dst_sorted, int* dst_orig, int* src_ptr, int* src_idx,
                            int* scratch, void* temp, long long temp_bytes, int* status, void* stream);

/* bf16 mode split of the same MLP: the hidden layers stay SIMT kernels, the 16 -> NO layer and its backward run as
 * K=16 / N=16 GEMMs of efl_tc_gemm_nt / efl_tc_gemm_tn.  hidden_fwd: x -> a2 [E,16] bf16 (input of the last Linear).
 * hidden_bwd: d a2 [E,16] fp32 -> dx [E,K0] and part [nblocks, efl_radial_hidden_part_size(K0)]:
 *   dW2 [16,16] | db2 | dg2 | dbeta2 | dW1 [16,K0] | db1 | dg1 | dbeta1. */
int efl_radial_hidden_part_size(int K0);
int efl_radial_hidden_fwd(const float* x, int K0, const float* W1, const float* b1, const float* g1, const float* be1,
                          const float* W2, const float* b2, const float* g2, const float* be2, void* a2_out, int E,
                          void* stream);
int efl_radial_hidden_bwd(const float* x, int K0, const float* W1, const float* b1, const float* g1, const float* be1,
                          const float* W2, const float* b2, const float* g2, const float* be2, const float* da2,
                          float* dx, float* part, int nblocks, int E, void* stream);

/* Deterministic gradient of a small-table embedding lookup (source/target element embeddings of the edge scalars,
 * Embedding.py:97-99): dW[r,c] = sum over lookups e with ids[e] = r of dx[e*ld + c], fixed summation order, no float
 * atomics.  ids [E] int64; dx already offset to the embedding's columns; workspace nblocks*rows*C floats. */
int efl_embedding_grad(const long long* ids, const float* dx, long long ld, int E, int rows, int C, float* dW,
                       float* workspace, int nblocks, void* stream);

/* EquivariantRMSNormArraySphericalHarmonicsV2 as built by get_normalization_layer (:2099-2192, :2254-2273):
 * centred l=0, degree-balanced RMS, per-degree gain gamma [lmax+1,16], bias beta [16] on l=0.  One warp per node.
 * bwd writes per-block partials of (d gamma | d beta) into part [nblocks, (lmax+2)*16]. */
int efl_rmsnorm_sh_fwd(int lmax, int C, const float* x, const float* gamma, const float* beta, float eps, float* y,
                       int N, void* stream);
int efl_rmsnorm_sh_bwd(int lmax, int C, const float* x, const float* gamma, float eps, const float* dy, float* dx,
                       float* part, int nblocks, int N, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EFL_H_ */
