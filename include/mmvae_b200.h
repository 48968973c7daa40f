/* mmvae_b200.h — C ABI of libmmvae_b200.so: the B200 (sm_100a) kernels behind bivae's multimodal-ELBO hot path.
 *
 * The reference (clementchadebec/mmvae, package `bivae`) has no FFI: its "plugin boundary" for this path is a
 * duck-typed Python API.  Each entry point below replaces one group of eager-PyTorch ops the
 * reference issues; the reference file:line it replaces is cited beside it (paths relative to
 * /root/reference/src/bivae).  The ctypes binding a maintainer would add is shown in INTEGRATION.md and is what
 * mmvae_b200/_lib.py does.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer to contiguous fp32 unless stated; nothing is owned or freed here;
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); all work is enqueued on it, no call
 *    synchronises, so every call is CUDA-graph capturable;
 *  - return value: 0 = ok, 1 = invalid argument, 1000+e = cudaError_t e.  mmvae_last_error() gives a message;
 *  - layouts: K-major sample stacking as in the reference: element (k,b) of a (K,B,...) tensor is row k*B+b
 *    (mmvae.py:41-54); evaluators use (B,K,...) (mmvae.py:133).
 */
#ifndef MMVAE_B200_H
#define MMVAE_B200_H
#ifdef __cplusplus
extern "C" {
#endif

#define MMVAE_ABI_VERSION 1

/* posterior / prior family of the latent (params.dist; multi_vaes.py:25) */
#define MMVAE_DIST_NORMAL   0
#define MMVAE_DIST_LAPLACE  1
/* decoder likelihood (params.recon_losses; multi_vaes.py:45-46) */
#define MMVAE_LIK_NORMAL    0
#define MMVAE_LIK_LAPLACE   1
#define MMVAE_LIK_BERNOULLI 2
/* "negative reconstruction loss" rows of m_jmvae_nf (recon_loss_dict, objectives.py:177,196-204): the same reduction
 * kernel with -F.mse_loss (no 1/2, no constant), -F.l1_loss, -F.binary_cross_entropy (logs clamped at -100) */
#define MMVAE_LIK_NEG_MSE   3
#define MMVAE_LIK_NEG_L1    4
#define MMVAE_LIK_NEG_BCE   5
/* how the second encoder head becomes a scale */
#define MMVAE_XFORM_EXP_HALF      0  /* std = exp(0.5*raw)            vae_model_adapted.py:84 */
#define MMVAE_XFORM_SOFTMAX       1  /* std = softmax(raw)*L + 1e-6   laplace_vae.py:69 */
#define MMVAE_XFORM_EXP_HALF_EPS  2  /* std = exp(0.5*raw) + 1e-6     joint_encoders.py:81-82 */
/* objective modes of the mixture-of-experts log-weight kernel */
#define MMVAE_OBJ_IWAE         0  /* objectives.py:117-131,333-340 */
#define MMVAE_OBJ_IWAE_LOOSER  1  /* objectives.py:343-369 */
#define MMVAE_OBJ_DREG         2  /* objectives.py:372-403 */
#define MMVAE_OBJ_DREG_LOOSER  3  /* objectives.py:406-438 */
/* GEMM epilogue activations */
#define MMVAE_ACT_NONE    0
#define MMVAE_ACT_RELU    1
#define MMVAE_ACT_SIGMOID 2

int         mmvae_abi_version(void);
const char* mmvae_last_error(void);
/* 1 if the library was compiled with the tcgen05/TMA GEMM and the device is sm_100; 0 otherwise */
int         mmvae_has_tcgen05(void);

/* ---- kernel (1): fused reparameterised rsample, all modalities and all K in one launch ---------------------
 * replaces laplace_vae.py:69-71, vae_model_adapted.py:84-86,133-137 on the K-tiled batch of mmvae.py:41-47.
 * mu,raw: [M,B,L]; noise: [M,K,B,L] (u~U(eps-1,1) for Laplace, eps~N(0,1) for Normal); out z: [M,K,B,L],
 * std: [M,B,L].  If bk_layout != 0 noise and z are [M,B,K,L] (evaluator layout). */
int mmvae_rsample_fwd(const float* mu, const float* raw, const float* noise, float* z, float* std,
                      int M, int K, int B, int L, int dist, int xform, int bk_layout, void* stream);
/* grad_z: [M,K,B,L]; grad_std_direct: [M,B,L] or NULL (gradient reaching std other than through z);
 * out grad_mu, grad_raw: [M,B,L]. */
int mmvae_rsample_bwd(const float* grad_z, const float* grad_std_direct, const float* noise, const float* raw,
                      const float* std, float* grad_mu, float* grad_raw,
                      int M, int K, int B, int L, int dist, int xform, int bk_layout, void* stream);

/* ---- kernel (2): decoder log-likelihood reduced over pixels ------------------------------------------------
 * replaces px_z(recon, 1).log_prob(x).view(K,B,-1).mul(lik_scaling).sum(-1) (objectives.py:381-385,416-419;
 * mmvae.py:154-156).  recon: [R,D]; x: [B,D]; row r uses x row (r / x_div) % B  (training: x_div=1; evaluator
 * (B,K) layout: x_div=K).  out lp: [R].  D % 4 == 0.  The row sum is carried in fp64; lp = its fp32 rounding and, if
 * lp_lo != NULL, lp_lo[R] = the residual (sum - lp) so that (lp, lp_lo) hands the softmax over K the digits an fp32
 * |lp| ~ 3000..9000 drops. */
int mmvae_loglik_fwd(const float* recon, const float* x, float* lp, float* lp_lo, int R, int B, int D, int x_div,
                     int lik, float lik_scaling, void* stream);
/* grad_lp: [R]; out grad: [R,D].  If through_sigmoid != 0, `recon` is sigmoid(h) and the output is the gradient
 * w.r.t. the pre-activation h (fuses the decoder's final Sigmoid backward, encoders.py:128, pythae Decoder_AE_MLP). */
int mmvae_loglik_bwd(const float* grad_lp, const float* recon, const float* x, float* grad, int R, int B, int D,
                     int x_div, int lik, float lik_scaling, int through_sigmoid, void* stream);

/* ---- kernel (3): cross-modal log q mixture, two-level logsumexp, objective and stop-gradient weights --------
 * replaces objectives.py:117-131 / 343-359 / 372-403 / 406-438 after model(x,K).
 * z: [M,K,B,L]; mu,std: [M,B,L]; lpx: [M(expert),M(decoder),K,B] (already lik-scaled), lpx_lo: its residuals from
 * mmvae_loglik_fwd or NULL; pz_loc,pz_scale: [L].  lw and the softmax over samples are formed in fp64.
 * out lw: [M,K,B]; wt: [M,K,B] = d loss / d lw (softmax weights incl. the 1/M of `looser`);
 * loss: [1] = objective (sum over datapoints); loss_per: [B] = each datapoint's term. */
int mmvae_moe_logw_fwd(const float* z, const float* mu, const float* std, const float* lpx, const float* lpx_lo,
                       const float* pz_loc, const float* pz_scale, float* lw, float* wt, float* loss,
                       float* loss_per, int M, int K, int B, int L, int dist, int mode, void* stream);
/* grad_loss: [1] (device scalar).  out grad_z [M,K,B,L], grad_lpx [M,M,K,B]; grad_mu, grad_std [M,B,L] are the
 * gradients through log q only (zero in the DReG modes: params detached, objectives.py:376,411). */
int mmvae_moe_logw_bwd(const float* grad_loss, const float* z, const float* mu, const float* std, const float* wt,
                       const float* pz_loc, const float* pz_scale, float* grad_z, float* grad_mu, float* grad_std,
                       float* grad_lpx, int M, int K, int B, int L, int dist, int mode, void* stream);

/* m_elbo (objectives.py:87-111): closed-form KL + importance ratio on the diagonal decoders.
 * lpx_diag: [M,K,B]; out loss [1], loss_per [B] (as above), terms [M*M] (sum over k,b of w*lpx for (r,d), the `details` dict). */
int mmvae_elbo_fwd(const float* z, const float* mu, const float* std, const float* lpx_diag,
                   const float* pz_loc, const float* pz_scale, float* loss, float* loss_per, float* terms,
                   int M, int K, int B, int L, int dist, void* stream);
int mmvae_elbo_bwd(const float* grad_loss, const float* z, const float* mu, const float* std,
                   const float* lpx_diag, const float* pz_loc, const float* pz_scale,
                   float* grad_mu, float* grad_std, float* grad_lpx_diag,
                   int M, int K, int B, int L, int dist, void* stream);

/* importance-sampled log-likelihood combine (mmvae.py:139-175, multi_vaes.py:290-316, jmvae_nf.py:231-268):
 * a: [B,Kc] log-weights of one K-shard -> per-datapoint running (max, sum exp(a-max)): mx,se: [B].
 * If accumulate != 0 the shard is merged into the existing (mx,se).  finish: ll[b] = mx + log(se) - log(K_total). */
int mmvae_lse_partial(const float* a, float* mx, float* se, int B, int Kc, int accumulate, void* stream);
int mmvae_lse_finish(const float* mx, const float* se, float* ll, int B, float log_K_total, void* stream);
/* evaluator log-weights: a[b,k] = sum_d lpx[d,b,k] + log p(z) - mixture term.  z: [B,K,L]; mu,std: [E,B,L]
 * (E experts in the proposal); qmode 0: -logsumexp_e(lq)/2 (mmvae.py:166 quirk), 1: -(logsumexp_e - log E),
 * 2: -lq_0 (single proposal), 3: no q term (z from the prior), 4: minus the given lq_ext[B,K]. */
int mmvae_eval_logw(const float* z, const float* mu, const float* std, const float* lpx, const float* lq_ext,
                    const float* pz_loc, const float* pz_scale, float* a,
                    int E, int D, int B, int K, int L, int dist, int qmode, void* stream);

/* ---- kernel (4): dense contractions ------------------------------------------------------------------------
 * C[M,N] = act( A * B + bias ), fp32 in/out, fp32 accumulate.  a_kmajor: A stored [M,K] (k contiguous) else
 * [K,M]; b_kmajor: B stored [N,K] (k contiguous, nn.Linear weight layout) else [K,N].  bias: [N] or NULL.
 * workspace: NULL or >= mmvae_gemm_workspace_bytes() bytes (split-K partials).  replaces torch.nn.Linear /
 * F.linear calls of pythae Encoder_VAE_MLP/Decoder_AE_MLP, joint_encoders.py:61-82, encoders.py:181-210 and,
 * through the lowering below, Conv2d/ConvTranspose2d of encoders.py:80-136; and their autograd backward. */
int mmvae_gemm(const float* A, const float* B, const float* bias, float* C, int M, int N, int K,
               int a_kmajor, int b_kmajor, int act, float* workspace, long workspace_bytes, int allow_tc,
               void* stream);
long mmvae_gemm_workspace_bytes(int M, int N, int K, int a_kmajor, int b_kmajor);
/* debug / speed probe of the tensor-core path: 0 = 3xTF32 split (fp32 accuracy, default), 1 = single TF32 pass.
 * Returns the previous mode. */
int mmvae_gemm_tc_set_mode(int mode);
/* number of tcgen05 GEMM launches (gemm_tc_kernel) this process has made: lets a test assert which kernel a shape ran on */
long mmvae_gemm_tc_launch_count(void);
/* which kernel the most recent mmvae_gemm call of this process ran on: 0 gemm_simt_kernel (tiled SIMT, exact fp32),
 * 1 gemm_tc_kernel (tcgen05), 2 gemm_skinny_kernel (K <= 32), 3 gemm_small_kernel (cluster split-K); -1 before the first call.
 * Accounting hook of bench.py / the tests. */
int mmvae_gemm_last_kernel(void);
/* contractions of at most this many multiply-adds (M*N*K) run on the one-launch cluster split-K kernel (gemm_small_kernel:
 * K split over the CTAs of a thread-block cluster, partials reduced through distributed shared memory, exact fp32);
 * 0 switches it off.  Returns the previous limit.  Tuning / test hook. */
long mmvae_gemm_small_set_limit(long macs);
/* column sums: out[n] = sum_m A[m,n]  (bias gradients); workspace >= mmvae_colsum_workspace_bytes(M,N) */
int  mmvae_colsum(const float* A, float* out, float* workspace, int M, int N, void* stream);
long mmvae_colsum_workspace_bytes(int M, int N);
/* mmvae_colsum without a workspace argument and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  The stage-1 partials live in a file-scope `static __device__` array of
 * mmvae_colsum_static_max_bytes() = 1 MiB that every GPU loading the library holds until exit; its device address is
 * resolved once per device (cudaGetSymbolAddress, with the caller's current device) and the call itself performs no
 * cudaMalloc / cudaFree / allocating library call.  Shapes whose partials (slabs(M,N) * N floats, the same count as
 * mmvae_colsum_workspace_bytes) exceed the array are rejected with code 1 and a message naming both sizes.  All calls
 * share the array: calls that may overlap on the GPU (different streams, no ordering) race on it; calls on one stream
 * are safe.  Nothing needs clearing between calls (every partial read is written first).  Result is bit-identical to
 * mmvae_colsum. */
int  mmvae_colsum_static(const float* A, float* out, int M, int N, void* stream);
long mmvae_colsum_static_max_bytes(void);
/* activation backward: out = g * act'(y)  (relu: y>0; sigmoid: y(1-y)); out may alias g */
int mmvae_act_bwd(const float* g, const float* y, float* out, long n, int act, void* stream);

/* conv lowering.  Activations are NHWC internally; generic element strides (sn,sc,sy,sx) describe the tensor at
 * the boundary.  Weights are used in permuted form: conv Wp[Cout,(ky,kx,ci)], convT Wp[Cin,(ky,kx,co)]. */
int mmvae_permute_weight(const float* w, float* wp, int C0, int C1, int kk, int to_permuted, void* stream);
/* im2col: x (N,C,H,W by strides) -> col[N*Ho*Wo, k*k*C]; Ho=(H+2p-k)/s+1 */
int mmvae_im2col(const float* x, float* col, int N, int C, int H, int W, int k, int s, int p,
                 long sn, long sc, long sy, long sx, void* stream);
/* col2im (gather form, deterministic): col[N*H*W, k*k*C] -> y (N,C,Ho,Wo by strides), Ho=(H-1)s-2p+k; adds
 * bias[C] (or NULL) and applies act. */
int mmvae_col2im(const float* col, const float* bias, float* y, int N, int C, int H, int W, int k, int s, int p,
                 long sn, long sc, long sy, long sx, int act, void* stream);

/* ConvTranspose2d(Cin -> Cout <= 4, k=4, s=2, p=1) as direct kernels (the image-producing layer of Decoder_VAE_SVHN,
 * encoders.py:125-126).  x: (N,H,W,Cin) NHWC; w: torch layout (Cin,Cout,4,4); y, g: (N,Cout,2H,2W) NCHW.
 * supported(): 0 = no, 1 = fwd + dgrad, 2 = fwd + dgrad + wgrad (Cin == 32). */
int  mmvae_convt_small_supported(int H, int W, int Cin, int Cout);
int  mmvae_convt_small_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin,
                           int Cout, int act, void* stream);
/* Evaluator form of the same forward (no gradient): the image is never written; lp[n] (+ lp_lo[n], the residual of the fp64
 * row sum, or NULL) = lik_scaling * sum_pix log p(xdata[(n / x_div) % B] | y_n) for the Normal / Laplace unit-scale likelihoods
 * (mmvae.py:154-156, multi_vaes.py:306-311 on the decoder output of encoders.py:125-128).  tcgen05 path only (H = W = 16,
 * Cin = 32): check _supported() first. */
int  mmvae_convt_small_fwd_loglik_supported(int H, int W, int Cin, int Cout, int lik);
int  mmvae_convt_small_fwd_loglik(const float* x, const float* w, const float* bias, const float* xdata, float* lp, float* lp_lo,
                                  int N, int H, int W, int Cin, int Cout, int act, int B, int x_div, int lik, float lik_scaling,
                                  void* stream);
/* g: gradient w.r.t. y; if y_act != NULL it is multiplied by act'(y_act) on the fly (g is then post-activation).
 * x_mask: NULL or the layer's input x (N,H,W,Cin): gx is zeroed where x <= 0 (fused ReLU backward of the layer below). */
int  mmvae_convt_small_dgrad(const float* g, const float* y_act, const float* w, const float* x_mask, float* gx,
                             int N, int H, int W, int Cin, int Cout, int act, void* stream);
/* gw: (Cin,Cout,4,4); gb: [Cout] or NULL; workspace >= mmvae_convt_small_wgrad_workspace_bytes(N,Cin,Cout). */
int  mmvae_convt_small_wgrad(const float* x, const float* g, const float* y_act, float* gw, float* gb, float* workspace,
                             int N, int H, int W, int Cin, int Cout, int act, void* stream);
long mmvae_convt_small_wgrad_workspace_bytes(int N, int Cin, int Cout);

/* ConvTranspose2d(k=4, s=2, p=1) with 32 | Cin, Cout <= 128 as an IMPLICIT GEMM on the tcgen05 path: the A operand is
 * fetched by 4-D TMA boxes over the NHWC tensors (shifted windows forward, stride-2 gathers backward; the padding ring
 * and ragged tiles are the TMA unit's out-of-bound zero fill), so no im2col / col2im buffer exists (encoders.py:119-124).
 * All tensors NHWC.  wf: class-major forward weights [4*Cout, 4*Cin] from mmvae_convt_weight_classes(w torch-layout);
 * wp / gwp: tap-major [Cin, 16*Cout] (mmvae_permute_weight layout).  g is the gradient w.r.t. the PRE-activation. */
int  mmvae_convt_tc_supported(int H, int W, int Cin, int Cout);
int  mmvae_convt_weight_classes(const float* w, float* wf, int Cin, int Cout, void* stream);
int  mmvae_convt_tc_fwd(const float* x, const float* wf, const float* bias, float* y, int Nimg, int H, int W, int Cin,
                        int Cout, int act, void* stream);
int  mmvae_convt_tc_dgrad(const float* g, const float* wp, const float* x_mask, float* gx, int Nimg, int H, int W,
                          int Cin, int Cout, void* stream);
int  mmvae_convt_tc_wgrad(const float* x, const float* g, float* gwp, float* workspace, long workspace_bytes, int Nimg,
                          int H, int W, int Cin, int Cout, void* stream);
long mmvae_convt_tc_wgrad_workspace_bytes(int Nimg, int H, int W, int Cin, int Cout);

/* Second generation of the same layer (32 | Cin, Cout; forward Cout in {32,64}; dgrad Cin in {32,64,128}; H*W | 128):
 * ConvTranspose2d(k4,s2,p1) forward and data gradient of encoders.py:113-124 with the activation operand read from TENSOR
 * MEMORY (tcgen05.mma "TS" form), operands split into bf16 (hi, mid) with hi*hi + hi*mid + mid*hi accumulated in fp32
 * (relative error <= 3*2^-18 per product), one TMA halo tile per 128-pixel tile instead of 9 / 16 shifted re-fetches, and
 * the weights resident in shared memory when they fit.  wf / wd: the layer's weights as bf16 operand matrices, written by
 * mmvae_convt_ts_pack_weights from the torch layout w (Cin,Cout,4,4); each holds mmvae_convt_ts_packed_elems() 16-bit
 * elements.  Tensors NHWC fp32 as above; g is the gradient w.r.t. the PRE-activation. */
int  mmvae_convt_ts_supported(int H, int W, int Cin, int Cout);
long mmvae_convt_ts_packed_elems(int Cin, int Cout);
int  mmvae_convt_ts_pack_weights(const float* w, void* wf, void* wd, int Cin, int Cout, void* stream);
int  mmvae_convt_ts_fwd(const float* x, const void* wf, const float* bias, float* y, int Nimg, int H, int W, int Cin,
                        int Cout, int act, void* stream);
int  mmvae_convt_ts_dgrad(const float* g, const void* wd, const float* x_mask, float* gx, int Nimg, int H, int W, int Cin,
                          int Cout, void* stream);
/* weight AND bias gradient of the same layer in one pass over x and g (Cin <= 128): gw in the torch layout (Cin,Cout,4,4),
 * gb [Cout] or NULL (replaces the autograd backward of ConvTranspose2d's weight / bias, encoders.py:113-124).  The pixel
 * contraction runs on tcgen05 with g transposed into tensor memory by the converter warps and x as an MN-major bf16 operand;
 * deterministic (partials in `workspace`, >= mmvae_convt_ts_wgrad_workspace_bytes(), added in a fixed order). */
long mmvae_convt_ts_wgrad_workspace_bytes(int Nimg, int H, int W, int Cin, int Cout);
int  mmvae_convt_ts_wgrad(const float* x, const float* g, float* gw, float* gb, float* workspace, long workspace_bytes,
                          int Nimg, int H, int W, int Cin, int Cout, void* stream);

/* MADE / MAF / IAF elementwise pieces (maf_model.py:62-74,91-104; iaf_model.py:60-78,94-100).  The masked-linear
 * contractions go through mmvae_gemm with the pre-masked weight (layers.py:24-25: F.linear(x, mask*W, b)). */
int mmvae_mask_weight(const float* w, const float* mask, float* out, long n, void* stream);
/* One MADE (made_model.py:21-111: masked Linear layers layers.py:14-25 with ReLU between them) on R rows as three kernels
 * instead of ~27 launches per training step -- the parallel direction of a flow block (MAF.forward maf_model.py:62-74,
 * IAF.inverse iaf_model.py:94-100) and its autograd backward.  dims: host array of n_layers+1 widths (input, hidden..., output;
 * multiples of 4, at most 128, equal hidden widths, 2 <= n_layers <= 5); ptrs: host array of 3*n_layers device pointers
 * {W_l [out,in], mask_l [out,in], bias_l [out] or NULL}; the masked weight mask*W is formed inside the kernels.
 * fwd: x [R,dims[0]] -> y [R,dims[n_layers]], h [n_layers-1][R][dims[1]] = the hidden activations (kept for backward).
 * bwd_data: gy -> gx (or NULL) and gpre [n_layers-1][R][dims[1]], the gradients w.r.t. the hidden pre-activations.
 * bwd_weights: grads = host array of 2*n_layers device pointers {gW_l, gb_l} (NULL = skip), each OVERWRITTEN with
 * mask_l * (gpre_l^T input_l) and the column sums of gpre_l (gpre of the last layer is gy); fixed summation order. */
int mmvae_made_supported(int n_layers, const int* dims);
int mmvae_made_fwd(const float* x, const void* ptrs, const int* dims, int n_layers, float* h, float* y, int R, void* stream);
int mmvae_made_bwd_data(const float* gy, const void* ptrs, const int* dims, int n_layers, const float* h, float* gpre,
                        float* gx, int R, void* stream);
int mmvae_made_bwd_weights(const float* x, const float* gy, const void* ptrs, const void* grads, const int* dims,
                           int n_layers, const float* h, const float* gpre, int R, void* stream);
/* The sequential direction of one MADE flow block (MAF.inverse maf_model.py:91-104; IAF.forward iaf_model.py:60-78; MADE
 * made_model.py:21-111) as ONE kernel: replaces L MADE passes (4 masked GEMMs each) + L affine steps.  Every hidden unit
 * is evaluated once, when the last column it depends on is produced.  y [R,L] block input (read at column L-1-i when
 * flip_in), out [R,L] (written at L-1-i when flip_out), ladj_out = ladj_in -/+ sum_i log_var_i (mode 0 / 1 as in
 * mmvae_flow_affine_fwd).  cnt[d] (d < L-1; host and device copies) = number of hidden units of degree <= d, at most 8
 * units per degree; pack = the degree-sorted masked weights, mmvae_made_seq_pack_floats(cnt, L) floats (layout:
 * csrc/made_seq.cu).  H = hidden width (three hidden layers).  Inference only (no backward). */
long mmvae_made_seq_pack_floats(const int* cnt_host, int L);
long mmvae_made_seq_smem_bytes(const int* cnt_host, int L, int H);
int mmvae_made_seq_fwd(const float* y, const float* ladj_in, float* out, float* ladj_out, const float* pack,
                       const int* cnt_dev, const int* cnt_host, int R, int L, int H, int mode, int flip_in, int flip_out,
                       void* stream);
/* affine coupling fused with the log|det J| update.  x: [R,L] block input; h: [R,2L] = MADE output (mu | log_var).
 * mode 0 "normalise" (MAF.forward, IAF.forward): u = (x-mu)*exp(-lv), ladj_out = ladj_in - sum lv;
 * mode 1 "generate"  (MAF.inverse, IAF.inverse): u = x*exp(lv)+mu,     ladj_out = ladj_in + sum lv.
 * col < 0: every column, out[r, flip_out ? L-1-l : l] = u_l (the flip of maf_model.py:74 folded in).
 * col >= 0: one step of the sequential direction: out = carry with column col replaced by u_col; ladj +-= lv[col]. */
int mmvae_flow_affine_fwd(const float* x, const float* h, const float* carry, const float* ladj_in, float* out,
                          float* ladj_out, int R, int L, int mode, int col, int flip_out, void* stream);
/* g_out [R,L], g_ladj [R] or NULL -> g_x [R,L], g_h [R,2L], g_carry [R,L] (col >= 0 only). */
int mmvae_flow_affine_bwd(const float* g_out, const float* g_ladj, const float* x, const float* h, float* g_x,
                          float* g_h, float* g_carry, int R, int L, int mode, int col, int flip_out, void* stream);
/* out[r, L-1-l] = x[r, l]  (x.flip(dims=(1,)), maf_model.py:74,91) */
int mmvae_flip_cols(const float* x, float* out, int R, int L, void* stream);
/* Gaussian row reductions of the JMVAE-NF regulariser.  z: [R,L]; mu,s: [R/div, L] (each shared by `div` rows).
 * kind 0: out[r] = sum_l log N(z; mu, std=s)                      (qz_xy.log_prob(z), jmvae_nf.py:69,260-261)
 * kind 1: out[r] = sum_l -0.5*(s + log 2pi + (z-mu)^2/exp(s)), s = log_var   (jmvae_nf.py:67,131-133)
 * kind 2: out[r] = -0.5*sum_l(1 + lv - mu^2 - exp(lv)), lv = 2 log s; z unused (objectives.py:208-211) */
int mmvae_gauss_rows_fwd(const float* z, const float* mu, const float* s, float* out, int R, int L, int kind,
                         int div, void* stream);
/* g: [R] -> g_z, g_mu, g_s: [R,L] (div == 1). */
int mmvae_gauss_rows_bwd(const float* g, const float* z, const float* mu, const float* s, float* g_z, float* g_mu,
                         float* g_s, int R, int L, int kind, void* stream);

/* fused multi-tensor Adam(amsgrad) over one flat buffer (main.py:114-115).  step_dev: device fp32 scalar holding
 * the number of steps taken so far; this call increments it first (so the update is CUDA-graph replayable). */
int mmvae_adam_amsgrad(float* p, const float* g, float* m, float* v, float* vmax, long n, float lr, float beta1,
                       float beta2, float eps, float* step_dev, int amsgrad, void* stream);

#ifdef __cplusplus
}
#endif
#endif
