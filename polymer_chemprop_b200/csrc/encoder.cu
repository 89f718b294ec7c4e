// C-ABI entry points and kernel sequencing of the wD-MPNN encoder (include/wdmpnn.h).
// Forward  == chemprop/models/mpn.py:92-171 (bond messages, dropout 0).
// Backward == autograd of the same, restated as deterministic CSR kernels.
#include <algorithm>
#include <array>
#include <cmath>
#include <map>
#include <mutex>

#include "common.cuh"
#include "graph_kernels.cuh"
#include "operands.cuh"
#include "simt_gemm.cuh"
#include "tc_host.cuh"

namespace wd {

thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
std::atomic<int64_t> g_graph_launches{0}, g_graph_builds{0};

namespace {

// ---- workspace layout (in floats; every section 256-byte aligned) --------------------------
struct Layout {
  int ldh;                 // row stride of every [*, h] activation buffer (h rounded up to 4)
  int64_t szB, szA;        // floats per [N_b, ldh] / [N_a, ldh] buffer
  int64_t I, M, Agg, H, mscale, U, dZ, dZ2, dS, dI, dA, dZo, partials, total;
  int64_t Po, Ph, Pi, Pcol;   // split-K partials of dW_o / dW_h (all steps) / dW_i, column-sum partials of db_o, db_h, db_i
  int64_t S;               // TF32 mode: S_t = A_t[b2a] - M_{t-1}[b2revb] written by the forward depth steps, read by dW_h
  int64_t Sign;            // TF32 mode + ReLU: sign bits of I and of every M_t (d x N_b x sign_ld words), see SignBits
  int sign_ld;             // words per row (over all sign tiles); 0 = the backward pass reads the activations themselves
  bool tc;                 // tensor-core math mode (TF32 or TF32X3)
  int Z;                   // max row chunks of the weight-gradient contractions (sizes `partials`)
  int Zh, Zi, Zo;          // row chunks of the dW_h / dW_i / dW_o contractions
  TcPacked pWi, pWh, pWo, pWhT, pWoT;   // TF32 mode: packed weight images (img == offset holder until bound)
  int64_t pk[5];           // offsets of the five packed images
  float* p(void* ws, int64_t off) const { return reinterpret_cast<float*>(ws) + off; }
};

constexpr int kColsumMaxChunks = 512;   // row chunks of the bias-gradient column sums

int dw_chunks(int64_t rows, int n1, int n2) {
  int64_t tiles = ceil_div(n1, DW_BM) * ceil_div(n2, DW_BN);
  int64_t z = ceil_div(4 * kNumSMs, tiles);
  z = std::min<int64_t>(z, std::max<int64_t>(1, rows / 256));
  return (int)std::max<int64_t>(1, std::min<int64_t>(z, 64));
}

// sign-bit geometry of a producer tiled like `B`: a sign tile is a column tile of the plain kernels (<= 160 columns)
// or an accumulator half of the gathered kernel (T4_HALF columns)
int sign_words(const TcPacked& B) {
  const int half = std::min(B.BN, T4_HALF);
  return B.n_tiles * (int)ceil_div(B.BN, half) * (int)ceil_div(half, 32);
}
// infer: forward-only layout (wd_encoder_infer) -- two alternating message buffers, one aggregate buffer, nothing
// kept for a backward pass (no S_t, no gradient scratch).
// Column-tile width of a tensor-core contraction over `rows` rows and N output columns.  Large batches take the widest
// tile the kernel allows (`cap`: every operand row is fetched once).  SMALL batches have fewer 128-row tiles than the
// GPU has SMs -- 26 at the reference's default batch of 50 graphs -- and a CTA's K loop is a serial chain of load ->
// transform -> MMA stages, so the kernel's duration is one CTA's latency: there the columns are split over more CTAs
// (down to 32 columns each) until ~one CTA per SM exists; the re-fetched operand rows are a few hundred KB from L2.
inline int adaptive_bn(int64_t rows, int N, int cap) {
  const int64_t m_tiles = ceil_div(std::max<int64_t>(rows, 1), TC_BM);
  const int np = (int)round_up(N, 16);
  int n_tiles = (int)ceil_div(np, cap);
  const int want = (int)std::min<int64_t>(ceil_div(kNumSMs, m_tiles), ceil_div(np, 32));
  n_tiles = std::max(n_tiles, want);
  return std::min(cap, (int)round_up(ceil_div(np, n_tiles), 16));
}

inline bool is_tc(const WdConfig& c) {
  return c.math_mode == WD_MATH_TF32 || c.math_mode == WD_MATH_TF32X3 || c.math_mode == WD_MATH_BF16X3;
}

Layout make_layout(const WdConfig& c, int64_t Na, int64_t Nb, int64_t G, bool infer = false) {
  Layout L;
  const int h = c.hidden_size;
  const bool tc = is_tc(c), bx3 = c.math_mode == WD_MATH_BF16X3, x3 = c.math_mode == WD_MATH_TF32X3 || bx3;
  L.tc = tc;
  L.ldh = (int)round_up(h, 8);   // rows start on 32-byte sectors: a 128-byte row chunk is exactly 4 sectors
  L.szB = round_up(Nb * L.ldh, 64);
  L.szA = round_up(Na * L.ldh, 64);
  int64_t off = 0;
  auto take = [&](int64_t n) { int64_t o = off; off += round_up(n, 64); return o; };
  L.I = take(L.szB);
  L.M = take((int64_t)(infer ? std::min(2, std::max(0, c.depth - 1)) : std::max(0, c.depth - 1)) * L.szB);
  L.Agg = take((int64_t)(infer ? 1 : c.depth) * L.szA);
  L.H = take(L.szA);
  L.mscale = take(G);
  L.U = take(c.undirected ? L.szB : 0);
  L.S = take(tc && !infer ? (int64_t)std::max(0, c.depth - 1) * L.szB : 0);
  L.dZ = take(infer ? 0 : L.szB);
  L.dZ2 = take(c.depth > 2 && !infer ? L.szB : 0);
  L.dS = take(infer ? 0 : L.szB);
  L.dI = take(infer ? 0 : L.szB);
  L.dA = take(infer ? 0 : L.szA);
  L.dZo = take(infer ? 0 : L.szA);
  const int Fo = c.atom_fdim + h;
  if (tc) {
    L.Zh = tc_dw_z(Nb, h, tc_dw_count_tiles(h, 0));
    L.Zi = tc_dw_z(Nb, h, tc_dw_count_tiles(c.bond_fdim, 0));
    L.Zo = tc_dw_z(Na, h, tc_dw_count_tiles(c.atom_fdim, h));
  } else {
    L.Zh = dw_chunks(Nb, h, h);
    L.Zi = dw_chunks(Nb, h, c.bond_fdim);
    L.Zo = dw_chunks(Na, h, Fo);
  }
  L.Z = std::max(std::max(L.Zh, L.Zi), L.Zo);
  // every reduction of the backward pass keeps its own partials until the single reduce launch at its end
  L.Po = take(infer ? 0 : (int64_t)L.Zo * h * Fo);
  L.Ph = take(infer ? 0 : (int64_t)std::max(0, c.depth - 1) * L.Zh * h * h);
  L.Pi = take(infer ? 0 : (int64_t)L.Zi * h * c.bond_fdim);
  L.Pcol = take(infer ? 0 : (int64_t)3 * kColsumMaxChunks * h);
  L.partials = L.Po;
  // forward weights: the 3xTF32 mode runs all three contractions on k_tc_gemm4 (tiles up to 304 columns, hi + lo
  // images); the single-pass mode keeps the plain ones on the TMA-fed k_tc_gemm3 (tiles up to 160 columns)
  // (3xBF16: the same kernel and tiles with ONE bf16 hi | lo image per chunk, tc_gemm4.cuh)
  const int fwd_bn = x3 ? T4_MAX_BN : T2_MAX_BN, parts = x3 && !bx3 ? 2 : 1, bf = bx3 ? 1 : 0;
  L.pWi = tc_packed_geometry(h, c.bond_fdim, 0, adaptive_bn(Nb, h, fwd_bn), parts, bf);     // I = f_bonds W_i^T
  L.pWh = tc_packed_geometry(h, h, 0, adaptive_bn(Nb, h, T4_MAX_BN), parts, bf);            // S W_h^T (gathered depth step)
  L.pWo = tc_packed_geometry(h, c.atom_fdim, h, adaptive_bn(Na, h, fwd_bn), parts, bf);     // [f_atoms, A_d] W_o^T
  L.pWhT = tc_packed_geometry(h, h, 0, adaptive_bn(Nb, h, T2_MAX_BN));   // dS = dZ W_h     (backward: single-pass TF32 in both modes)
  L.pWoT = tc_packed_geometry(h, h, 0, adaptive_bn(Na, h, T2_MAX_BN));   // dA_d = dZo W_o[:, F_a:]
  // the message gradient only needs (y > 0) of a ReLU output: the tensor-core epilogues leave those bits behind
  // (decided by the config alone, so that the forward and the backward call of one step cannot disagree)
  L.sign_ld = 0;
  if (tc && !infer && c.activation == WD_ACT_RELU && !(c.flags & WD_FLAG_NO_SIGN_BITS))
    L.sign_ld = std::max(sign_words(L.pWi), sign_words(L.pWh));
  L.Sign = take((int64_t)c.depth * Nb * L.sign_ld);
  const TcPacked* pk[5] = {&L.pWi, &L.pWh, &L.pWo, &L.pWhT, &L.pWoT};
  for (int i = 0; i < 5; ++i) L.pk[i] = take(tc && !(infer && i >= 3) ? pk[i]->floats() : 0);
  L.total = off;
  return L;
}

int check_cfg(const WdConfig* c) {
  if (!c) return fail(WD_ERR_INVALID_ARG, "cfg is NULL");
  if (c->hidden_size < 1 || c->atom_fdim < 1 || c->bond_fdim < 1 || c->depth < 1)
    return fail(WD_ERR_INVALID_ARG, "hidden_size/atom_fdim/bond_fdim/depth must be >= 1");
  if (c->activation < WD_ACT_RELU || c->activation > WD_ACT_ELU)
    return fail(WD_ERR_UNSUPPORTED, "activation %d not supported (ReLU, LeakyReLU, tanh, SELU, ELU)", c->activation);
  if (c->aggregation < WD_AGG_MEAN || c->aggregation > WD_AGG_NORM)
    return fail(WD_ERR_INVALID_ARG, "aggregation %d unknown", c->aggregation);
  if (c->math_mode != WD_MATH_FP32 && c->math_mode != WD_MATH_TF32 && c->math_mode != WD_MATH_TF32X3 &&
      c->math_mode != WD_MATH_BF16X3)
    return fail(WD_ERR_INVALID_ARG, "math_mode %d unknown", c->math_mode);
  if (!(c->dropout >= 0.f && c->dropout < 1.f))
    return fail(WD_ERR_INVALID_ARG, "dropout %g must be in [0, 1)", (double)c->dropout);
  return WD_OK;
}

int check_graph(const WdGraph* g, const WdWeights* w, const WdConfig* c) {
  if (!g || !w) return fail(WD_ERR_INVALID_ARG, "graph/weights is NULL");
  if (g->n_atoms < 1 || g->n_bonds < 1 || g->n_mols < 0)
    return fail(WD_ERR_INVALID_ARG, "graph sizes invalid (n_atoms=%lld n_bonds=%lld n_mols=%lld)",
                (long long)g->n_atoms, (long long)g->n_bonds, (long long)g->n_mols);
  if (!g->f_atoms || !g->f_bonds || !g->w_atoms || !g->w_bonds || !g->in_ptr || !g->src || !g->rev ||
      !g->dst || !g->mol_of || (g->n_bonds > 1 && (!g->in_idx || !g->out_idx)) ||
      (g->n_mols > 0 && (!g->a_start || !g->a_size || !g->dop)))
    return fail(WD_ERR_INVALID_ARG, "graph has NULL device pointers");
  if ((g->ld_atoms && g->ld_atoms < c->atom_fdim) || (g->ld_bonds && g->ld_bonds < c->bond_fdim))
    return fail(WD_ERR_INVALID_ARG, "ld_atoms / ld_bonds must be 0 or >= atom_fdim / bond_fdim");
  if (!w->W_i || !w->W_h || !w->W_o || !w->b_o || (c->bias && (!w->b_i || !w->b_h)))
    return fail(WD_ERR_INVALID_ARG, "weights have NULL device pointers");
  return WD_OK;
}

// dropout site `salt` (depth step t = 1 .. d-1, d = atom hiddens) of this call; thresh == 0 switches it off
Dropout make_dropout(const WdConfig& c, int salt) {
  Dropout d{};
  if (c.dropout > 0.f) {
    d.thresh = (uint32_t)std::min<double>(4294967295.0, std::floor((double)c.dropout * 4294967296.0));
    d.scale = 1.f / (1.f - c.dropout);
    d.seed_lo = c.dropout_seed_lo;
    d.seed_hi = c.dropout_seed_hi;
    d.salt = (uint32_t)salt;
  }
  return d;
}

// sign bits of activation `t` (0: input projection I, t >= 1: message M_t), produced by a kernel tiled like `B`
SignBits sign_bits(const Layout& L, void* ws, int64_t Nb, int t, const TcPacked& B) {
  if (L.sign_ld == 0) return SignBits{};
  const int half = std::min(B.BN, T4_HALF);
  return SignBits{reinterpret_cast<uint32_t*>(L.p(ws, L.Sign)) + (int64_t)t * Nb * L.sign_ld, Nb, B.BN, half,
                  (int)ceil_div(B.BN, half), (int)ceil_div(half, 32)};
}

// rows per block of the kernels that walk block-local runs of rows (k_message_grad, k_aggregate, k_readout_grad):
// up to MG_ROWS, but at least ~16 blocks per SM so that small batches still fill the machine
inline int rows_per_block(int64_t n_rows) {
  return (int)std::max<int64_t>(4, std::min<int64_t>(MG_ROWS, n_rows / (16 * kNumSMs)));
}

inline int grid1d(int64_t n, int block = 256) { return (int)ceil_div(std::max<int64_t>(n, 1), block); }

// ---- contraction dispatch: FP32 -> SIMT FFMA kernels, TF32 -> tcgen05 kernels (tc == packed weights) ----
template <int AMODE, int BMODE, int EPI, int ACT, bool PREACT>
int gemm(const WdConfig& c, cudaStream_t s, const AOperand& A, const float* W, int ldb, const float* bias,
         const float* R, int ldr, float* C, int ldc, int64_t M, int N, int K, const TcPacked* tc,
         const Dropout& drop = Dropout{}, const SignBits& sign = SignBits{}) {
  if (is_tc(c)) return tc_gemm<AMODE, EPI, ACT, PREACT>(s, A, *tc, bias, R, ldr, C, ldc, M, drop, sign);
  dim3 grid((unsigned)ceil_div(M, SG_BM), (unsigned)ceil_div(std::max(N, ldc), SG_BN));
  WD_LAUNCH((k_simt_gemm<AMODE, BMODE, EPI, ACT, PREACT>), grid, SG_THREADS, 0, s, A, W, ldb, bias, R, ldr,
            C, ldc, M, N, K, drop);
  return WD_OK;
}

// P[0 .. Z) <- partial X^T Y over Z row chunks; Y has w0 (+ w1 for CONCAT) columns
template <int YMODE, int ACT, bool PREACT>
int dw_partial(const WdConfig& c, cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1,
               float* P, int64_t R, int N1, int Z) {
  if (is_tc(c)) {
    if constexpr (YMODE == A_GATHER_DIFF) return fail(WD_ERR_UNSUPPORTED, "tensor-core dW reads S_t, not the gathered operand");
    else return tc_dw<YMODE>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  }
  const int N2 = w0 + w1;
  int64_t rpc = round_up(ceil_div(R, Z), DW_BK);
  dim3 grid((unsigned)ceil_div(N2, DW_BN), (unsigned)ceil_div(N1, DW_BM), (unsigned)Z);
  WD_LAUNCH((k_simt_dw<YMODE, ACT, PREACT>), grid, DW_THREADS, 0, s, X, ldx, Y, P, R, N1, N2, rpc);
  return WD_OK;
}

// TF32 mode: bind the packed-weight images to the workspace and (re)pack them from the live weights
int bind_packed(const WdConfig& c, Layout& L, void* ws) {
  TcPacked* pk[5] = {&L.pWi, &L.pWh, &L.pWo, &L.pWhT, &L.pWoT};
  for (int i = 0; i < 5; ++i) pk[i]->img = is_tc(c) ? L.p(ws, L.pk[i]) : nullptr;
  return WD_OK;
}

// Reductions of one pass, collected and issued as ONE launch (ReduceJobs); every job keeps its own partial buffer
// until then.  The order of the additions inside a job is fixed, so the result does not depend on the batching.
struct Reducer {
  ReduceJobs jobs{};
  bool vec = true;        // every job so far has n % 4 == 0 and 16-byte aligned partials: four outputs per thread
  int add(const float* P, float* out, int64_t n, int Z) {
    if (jobs.count >= RP_MAX_JOBS) return fail(WD_ERR_UNSUPPORTED, "too many reductions in one pass");
    jobs.j[jobs.count++] = ReduceJob{P, out, n, Z, 0u};
    vec = vec && n % 4 == 0 && (reinterpret_cast<uintptr_t>(P) & 15) == 0;
    return WD_OK;
  }
  int run(cudaStream_t s) {
    if (jobs.count == 0) return WD_OK;
    ProfScope ps(WD_PROF_REDUCE);
    const int per_block = RP_COLS * (vec ? 4 : 1);
    unsigned blocks = 0;
    for (int k = 0; k < jobs.count; ++k) {
      jobs.j[k].block0 = blocks;
      blocks += (unsigned)ceil_div(jobs.j[k].n, per_block);
    }
    if (vec) WD_LAUNCH(k_reduce_partials<4>, blocks, RP_COLS * RP_GROUPS, 0, s, jobs);
    else     WD_LAUNCH(k_reduce_partials<1>, blocks, RP_COLS * RP_GROUPS, 0, s, jobs);
    jobs.count = 0;
    vec = true;
    return WD_OK;
  }
};

int reduce_partials(cudaStream_t s, const float* P, float* out, int64_t n, int Z) {
  Reducer r;
  r.add(P, out, n, Z);
  return r.run(s);
}

// bias gradients: column sums over many short row chunks (coalesced row reads) into P; the fixed-order reduction of
// the chunks is queued on `red`
int colsum_chunks(int64_t R) { return (int)std::max<int64_t>(1, std::min<int64_t>(kColsumMaxChunks, R / 64)); }
int colsum(cudaStream_t s, const float* X, int ldx, float* P, float* out, int64_t R, int N, Reducer& red) {
  const int Z = colsum_chunks(R);
  int64_t rpc = ceil_div(R, Z);
  { ProfScope ps(WD_PROF_REDUCE);
  if (N % 4 == 0 && ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0)
    WD_LAUNCH(k_colsum_partial4, dim3((unsigned)ceil_div(N, 4 * 96), (unsigned)Z), 96, 0, s, X, ldx, P, R, N, rpc);
  else
    WD_LAUNCH(k_colsum_partial, dim3((unsigned)ceil_div(N, 128), (unsigned)Z), 128, 0, s, X, ldx, P, R, N, rpc); }
  return red.add(P, out, N, Z);
}

#define WD_TRY(expr) do { int _rc = (expr); if (_rc != WD_OK) return _rc; } while (0)

// out[a] = sum over the in-edge (incoming == true, weighted by w) or out-edge list of atom a of U(b)
template <int ACT, bool PREACT>
int aggregate(cudaStream_t s, const WdGraph& g, bool incoming, const float* U, int ld, const float* w, float* out) {
  const int nc4 = ld / 4;
  const int64_t Na = g.n_atoms;
  // (a padded [N_a, K] int32 form of the lists -- one 16-byte id load at a fixed address instead of the ptr -> idx
  //  chain -- was measured 5 % SLOWER on B200, profiles/r1_j_padded_lists_vs_csr.txt, and removed)
  const int rpb = rows_per_block(Na);
  WD_LAUNCH((k_aggregate<ACT, PREACT>), (unsigned)ceil_div(Na, rpb), 256, 0, s, U, ld, w, g.in_ptr,
            incoming ? g.in_idx : g.out_idx, out, Na, nc4, rpb);
  return WD_OK;
}

// ---- forward ----------------------------------------------------------------------------------
template <int ACT>
int forward_impl(const WdGraph& g, const WdWeights& w, const WdConfig& c, void* ws, float* out,
                 cudaStream_t s, bool infer) {
  Layout L = make_layout(c, g.n_atoms, g.n_bonds, g.n_mols, infer);
  const int h = c.hidden_size, ld = L.ldh, nc4 = ld / 4, Fa = c.atom_fdim, Fb = c.bond_fdim;
  const int64_t Na = g.n_atoms, Nb = g.n_bonds, G = g.n_mols;
  float* I = L.p(ws, L.I);
  bind_packed(c, L, ws);
  if (L.tc) {   // weights change every step: repack (tiny) on the stream
    ProfScope ps(WD_PROF_MISC);
    TcPacker pk;                       // one launch for the three forward weights
    pk.add(w.W_i, Fb, B_NK, L.pWi);
    if (c.depth > 1) pk.add(w.W_h, h, B_NK, L.pWh);
    pk.add(w.W_o, Fa + h, B_NK, L.pWo);
    WD_TRY(pk.run(s));
  }
  // input = W_i(f_bonds)                                                      mpn.py:96
  AOperand a_in{g.f_bonds, nullptr, nullptr, nullptr, g.ld_bonds ? g.ld_bonds : Fb, 0, 0};
  { ProfScope ps(WD_PROF_INPUT_PROJ);
  WD_TRY((gemm<A_PLAIN, B_NK, E_STORE, ACT, false>(c, s, a_in, w.W_i, Fb, c.bias ? w.b_i : nullptr, nullptr, 0,
                                                  I, ld, Nb, h, Fb, &L.pWi, Dropout{}, sign_bits(L, ws, Nb, 0, L.pWi)))); }
  const float* prev = I;   // message = act(input) is applied on load (PREACT)          :97
  for (int t = 1; t < c.depth; ++t) {                                               // :100
    const bool pre = (t == 1);
    float* A_t = L.p(ws, L.Agg) + (int64_t)(infer ? 0 : t - 1) * L.szA;
    float* M_t = L.p(ws, L.M) + (int64_t)(infer ? (t - 1) & 1 : t - 1) * L.szB;
    const float* u = prev;
    bool upre = pre;
    if (c.undirected) {                                                             // :101-102
      float* U = L.p(ws, L.U);
      if (pre) WD_LAUNCH((k_symmetrize<ACT, true>), grid1d(Nb * nc4), 256, 0, s, prev, ld, g.rev, U, Nb, nc4);
      else     WD_LAUNCH((k_symmetrize<ACT, false>), grid1d(Nb * nc4), 256, 0, s, prev, ld, g.rev, U, Nb, nc4);
      u = U;
      upre = false;
    }
    // a_message = sum_k w_bonds[a2b] * message[a2b]                                  :112-118
    { ProfScope ps(WD_PROF_AGGREGATE);
    if (upre) WD_TRY((aggregate<ACT, true>(s, g, true, u, ld, g.w_bonds, A_t)));
    else      WD_TRY((aggregate<ACT, false>(s, g, true, u, ld, g.w_bonds, A_t))); }
    ProfScope ps(WD_PROF_DEPTH_STEP);
    // message = act(input + W_h(a_message[b2a] - message[b2revb]))                   :119-123
    AOperand a_s{A_t, u, g.src, g.rev, ld, ld, 0};
    a_s.rev_adjacent = g.rev_adjacent;
    if (L.tc && !infer) a_s.S_out = L.p(ws, L.S) + (int64_t)(t - 1) * L.szB;   // kept for dW_h
    const float* bh = c.bias ? w.b_h : nullptr;
    const Dropout drop = make_dropout(c, t);                                          // :124
    const SignBits sg = sign_bits(L, ws, Nb, t, L.pWh);
    if (upre) WD_TRY((gemm<A_GATHER_DIFF, B_NK, E_RES_ACT, ACT, true>(c, s, a_s, w.W_h, h, bh, I, ld, M_t, ld, Nb, h, h, &L.pWh, drop, sg)));
    else      WD_TRY((gemm<A_GATHER_DIFF, B_NK, E_RES_ACT, ACT, false>(c, s, a_s, w.W_h, h, bh, I, ld, M_t, ld, Nb, h, h, &L.pWh, drop, sg)));
    prev = M_t;
  }
  // final aggregate + atom update                                                    :127-134
  float* A_d = L.p(ws, L.Agg) + (int64_t)(infer ? 0 : c.depth - 1) * L.szA;
  { ProfScope ps(WD_PROF_AGGREGATE);
  if (c.depth == 1) WD_TRY((aggregate<ACT, true>(s, g, true, prev, ld, g.w_bonds, A_d)));
  else              WD_TRY((aggregate<ACT, false>(s, g, true, prev, ld, g.w_bonds, A_d))); }
  float* H = L.p(ws, L.H);
  AOperand a_o{g.f_atoms, A_d, nullptr, nullptr, g.ld_atoms ? g.ld_atoms : Fa, ld, Fa};
  { ProfScope ps(WD_PROF_ATOM_UPDATE);
  WD_TRY((gemm<A_CONCAT, B_NK, E_ACT, ACT, false>(c, s, a_o, w.W_o, Fa + h, w.b_o, nullptr, 0, H, ld, Na, h, Fa + h, &L.pWo,
                                                  make_dropout(c, c.depth)))); }                 // :133-134
  // readout                                                                          :146-171
  if (G > 0) {
    ProfScope ps(WD_PROF_READOUT);
    float* ms = L.p(ws, L.mscale);
    WD_LAUNCH(k_mol_scale, grid1d(G), 256, 0, s, g.w_atoms, g.a_start, g.a_size, g.dop, c.aggregation,
              c.aggregation_norm, ms, G);
    WD_LAUNCH(k_readout, grid1d(G * h), 256, 0, s, H, ld, g.w_atoms, g.a_start, g.a_size, ms, out, h, G);
  }
  return WD_OK;
}

// ---- backward ---------------------------------------------------------------------------------
template <int ACT, bool PREACT, bool HAS_DS, bool WRITE_DZ, bool WRITE_DI>
int message_grad(const WdConfig& c, cudaStream_t s, const WdGraph& g, const float* dA, const float* dS,
                 const float* Y, int ld, float* dZ, const float* dI_in, float* dI, const Dropout& drop,
                 const SignBits& sb, const float* dI_in2 = nullptr) {
  const int nc4 = ld / 4;
  const int64_t Nb = g.n_bonds;
  ProfScope ps(WD_PROF_MSG_GRAD);
  constexpr bool kRelu = ACT == WD_ACT_RELU;   // the only activation whose derivative the sign bits carry
  const int rpb = rows_per_block(Nb);
  if (kRelu && sb.p != nullptr) {
    if (c.undirected && HAS_DS)
      WD_LAUNCH((k_message_grad<ACT, PREACT, HAS_DS, true, WRITE_DZ, WRITE_DI, kRelu>), (unsigned)ceil_div(Nb, rpb), 256, 0, s, dA, dS, Y,
                ld, g.w_bonds, g.dst, g.rev, dZ, dI_in, dI_in2, dI, Nb, nc4, drop, sb, rpb);
    else
      WD_LAUNCH((k_message_grad<ACT, PREACT, HAS_DS, false, WRITE_DZ, WRITE_DI, kRelu>), (unsigned)ceil_div(Nb, rpb), 256, 0, s, dA, dS, Y,
                ld, g.w_bonds, g.dst, g.rev, dZ, dI_in, dI_in2, dI, Nb, nc4, drop, sb, rpb);
    return WD_OK;
  }
  if (c.undirected && HAS_DS)
    WD_LAUNCH((k_message_grad<ACT, PREACT, HAS_DS, true, WRITE_DZ, WRITE_DI, false>), (unsigned)ceil_div(Nb, rpb), 256, 0, s, dA, dS, Y,
              ld, g.w_bonds, g.dst, g.rev, dZ, dI_in, dI_in2, dI, Nb, nc4, drop, SignBits{}, rpb);
  else
    WD_LAUNCH((k_message_grad<ACT, PREACT, HAS_DS, false, WRITE_DZ, WRITE_DI, false>), (unsigned)ceil_div(Nb, rpb), 256, 0, s, dA, dS, Y,
              ld, g.w_bonds, g.dst, g.rev, dZ, dI_in, dI_in2, dI, Nb, nc4, drop, SignBits{}, rpb);
  return WD_OK;
}

// A second branch of a captured pass (small batches, GraphRun below): work that is off the critical path -- the weight
// gradients, the bias column sums, the weight packing -- is captured on a side stream, so the executable graph may run it
// next to the dA -> dM -> dS -> dA chain instead of inside it.  At batch 50 no kernel fills the GPU for long (8-24 us
// each), the chain is latency-bound, and four of its sixteen links are weight-gradient kernels nothing downstream waits
// for until the final reduction.  Without a side stream (eager launches, large batches) every call is a no-op and the
// work stays on the main stream, in the same order.
struct SideBranch {
  cudaStream_t main = nullptr, side = nullptr;
  cudaEvent_t ev[8] = {};
  int next = 0;
  bool on() const { return side != nullptr; }
  cudaStream_t stream() const { return side ? side : main; }
  int fork() {          // the side branch waits for everything launched on the main stream so far
    if (!side) return WD_OK;
    cudaEvent_t e = ev[next++ & 7];
    WD_CUDA_OK(cudaEventRecord(e, main));
    WD_CUDA_OK(cudaStreamWaitEvent(side, e, 0));
    return WD_OK;
  }
  int join() {          // the main stream waits for everything launched on the side branch so far
    if (!side) return WD_OK;
    cudaEvent_t e = ev[next++ & 7];
    WD_CUDA_OK(cudaEventRecord(e, side));
    WD_CUDA_OK(cudaStreamWaitEvent(main, e, 0));
    pdl_guard().plain_once = true;     // the next main-stream kernel has two predecessors: full edges (common.cuh)
    return WD_OK;
  }
};

template <int ACT>
int backward_impl(const WdGraph& g, const WdWeights& w, const WdConfig& c, void* ws, const float* d_out,
                  const WdGrads& gr, cudaStream_t s, SideBranch sd = SideBranch{}) {
  sd.main = s;
  const cudaStream_t s2 = sd.stream();     // where the off-chain work goes: the side branch when there is one
  struct PlainSide {                       // launches on the side branch carry no programmatic-launch attribute (common.cuh)
    explicit PlainSide(cudaStream_t st) { pdl_guard().plain_stream = st; }
    ~PlainSide() { pdl_guard().plain_stream = nullptr; pdl_guard().plain_once = false; }
  } plain_side(sd.side);
  Layout L = make_layout(c, g.n_atoms, g.n_bonds, g.n_mols);
  const int h = c.hidden_size, ld = L.ldh, nc4 = ld / 4, Fa = c.atom_fdim, Fb = c.bond_fdim, Fo = Fa + h;
  bind_packed(c, L, ws);
  if (L.tc) {
    ProfScope ps(WD_PROF_MISC);
    TcPacker pk;                       // one launch for the two transposed backward weights
    if (c.depth > 1) pk.add(w.W_h, h, B_KN, L.pWhT);
    pk.add(w.W_o + Fa, Fo, B_KN, L.pWoT);
    WD_TRY(sd.fork());                 // (side branch: next to the readout gradient, joined before the first dX)
    WD_TRY(pk.run(s2));
  }
  const int64_t Na = g.n_atoms, Nb = g.n_bonds;
  const int d = c.depth;
  float *I = L.p(ws, L.I), *H = L.p(ws, L.H), *dS = L.p(ws, L.dS), *dI = L.p(ws, L.dI),
        *dA = L.p(ws, L.dA), *dZo = L.p(ws, L.dZo);
  float *Po = L.p(ws, L.Po), *Ph = L.p(ws, L.Ph), *Pi = L.p(ws, L.Pi), *Pcol = L.p(ws, L.Pcol);
  const int64_t pcol = (int64_t)kColsumMaxChunks * h;    // one column-sum region per bias gradient
  Reducer red;                                            // all reductions of this pass: one launch at the end
  float* A_d = L.p(ws, L.Agg) + (int64_t)(d - 1) * L.szA;

  // readout + atom-update activation:  dZo = w_a * dop/den * d_out[mol] * act'(H)
  { ProfScope ps(WD_PROF_READOUT);
  const int rpb = rows_per_block(Na);
  if (h % 4 == 0 && (reinterpret_cast<uintptr_t>(d_out) & 15) == 0)
    WD_LAUNCH((k_readout_grad<ACT, true>), (unsigned)ceil_div(Na, rpb), 256, 0, s, d_out, h, H, ld, g.w_atoms, g.mol_of,
              L.p(ws, L.mscale), dZo, Na, make_dropout(c, c.depth), rpb);
  else
    WD_LAUNCH((k_readout_grad<ACT, false>), (unsigned)ceil_div(Na, rpb), 256, 0, s, d_out, h, H, ld, g.w_atoms, g.mol_of,
              L.p(ws, L.mscale), dZo, Na, make_dropout(c, c.depth), rpb); }
  // dZo (114 MB at B=4096) has just been written and fits L2: its three readers run back to back, smallest first
  if (L.tc) WD_TRY(sd.join());         // the packed weights are ready for the main stream
  WD_TRY(sd.fork());                   // dZo is ready for the side branch
  {  // db_o = colsum(dZo)
    ProfScope ps(WD_PROF_DW);
    WD_TRY(colsum(s2, dZo, ld, Pcol, gr.db_o, Na, h, red));
  }
  // dA_d = dZo * W_o[:, F_a:]
  AOperand a_zo{dZo, nullptr, nullptr, nullptr, ld, 0, 0};
  { ProfScope ps(WD_PROF_DX);
  WD_TRY((gemm<A_PLAIN, B_KN, E_STORE, ACT, false>(c, s, a_zo, w.W_o + Fa, Fo, nullptr, nullptr, 0, dA, ld, Na, h, h, &L.pWoT))); }
  {  // dW_o = dZo^T [f_atoms, A_d]
    ProfScope ps(WD_PROF_DW);
    const int Z = L.Zo;
    AOperand y{g.f_atoms, A_d, nullptr, nullptr, g.ld_atoms ? g.ld_atoms : Fa, ld, Fa};
    WD_TRY((dw_partial<A_CONCAT, ACT, false>(c, s2, dZo, ld, y, Fa, h, Po, Na, h, Z)));
    WD_TRY(red.add(Po, gr.dW_o, (int64_t)h * Fo, Z));
  }

  const int Zh = L.Zh;
  // dZ_t alternates between two buffers so that step t-1 can read dZ_t (as the running sum) while it writes its own;
  // `sum` points at sum_{t' >= t} dZ_t': the first step's dZ itself (no separate dI write), dI afterwards
  float* dZbuf[2] = {L.p(ws, L.dZ), L.p(ws, L.dZ2)};
  const float* sum = nullptr;
  // depth 3 without a W_h bias: both dZ_2 and dZ_1 are still in their buffers when the last step runs, so the running
  // sum dI = dZ_2 + dZ_1 is not materialised by the middle step (one read and one write of [N_b, h] less per pass);
  // the last step forms (dZ_2 + dZ_1) + dZ_0 itself -- the same association, hence the same bits
  const bool defer_sum = d == 3 && !(c.bias && gr.db_h) && !(c.flags & WD_FLAG_NO_DEFER_SUM);
  const float* sum2 = nullptr;
  for (int t = d - 1; t >= 1; --t) {
    float* dZ = dZbuf[(d - 1 - t) & 1];
    const float* M_t = L.p(ws, L.M) + (int64_t)(t - 1) * L.szB;
    const float* A_t = L.p(ws, L.Agg) + (int64_t)(t - 1) * L.szA;
    const float* M_prev = (t == 1) ? I : L.p(ws, L.M) + (int64_t)(t - 2) * L.szB;
    const bool pre = (t == 1);
    // dZ_t = dM_t * act'(M_t), dI += dZ_t ; dM_t = w_b dA[dst b] - dS[rev b] (symmetrised if undirected)
    const SignBits sg = sign_bits(L, ws, Nb, t, L.pWh);
    if (d - 1 - t >= 2) WD_TRY(sd.join());   // this step overwrites the dZ buffer the weight gradient of two steps ago reads
    if (t == d - 1) { WD_TRY((message_grad<ACT, false, false, true, false>(c, s, g, dA, dS, M_t, ld, dZ, nullptr, dI, make_dropout(c, t), sg))); sum = dZ; }
    else if (defer_sum) { WD_TRY((message_grad<ACT, false, true, true, false>(c, s, g, dA, dS, M_t, ld, dZ, nullptr, dI, make_dropout(c, t), sg))); sum2 = dZ; }
    else            { WD_TRY((message_grad<ACT, false, true, true, true>(c, s, g, dA, dS, M_t, ld, dZ, sum, dI, make_dropout(c, t), sg))); sum = dI; }
    // dW_h += dZ_t^T S_t, S_t = A_t[b2a] - U_{t-1}[b2revb]: recomputed by the gather (FP32 path) or read back (TF32 path)
    const float* u = M_prev;
    bool upre = pre;
    const bool have_S = L.tc;   // forward kept S_t (tensor-core path)
    if (c.undirected && !have_S) {
      float* U = L.p(ws, L.U);
      WD_TRY(sd.join());               // the previous step's weight gradient (side branch) still reads U
      if (pre) WD_LAUNCH((k_symmetrize<ACT, true>), grid1d(Nb * nc4), 256, 0, s, M_prev, ld, g.rev, U, Nb, nc4);
      else     WD_LAUNCH((k_symmetrize<ACT, false>), grid1d(Nb * nc4), 256, 0, s, M_prev, ld, g.rev, U, Nb, nc4);
      u = U;
      upre = false;
    }
    AOperand y{A_t, u, g.src, g.rev, ld, ld, 0};
    float* Pt = Ph + (int64_t)(d - 1 - t) * Zh * h * h;
    WD_TRY(sd.fork());                 // dZ_t (and the symmetrised U) are ready for the side branch
    { ProfScope ps(WD_PROF_DW);
    if (have_S) {   // the forward pass left S_t in the workspace: a plain TMA-fed contraction
      AOperand ys{L.p(ws, L.S) + (int64_t)(t - 1) * L.szB, nullptr, nullptr, nullptr, ld, 0, 0};
      WD_TRY((dw_partial<A_PLAIN, ACT, false>(c, s2, dZ, ld, ys, h, 0, Pt, Nb, h, Zh)));
    } else if (upre) WD_TRY((dw_partial<A_GATHER_DIFF, ACT, true>(c, s2, dZ, ld, y, h, 0, Pt, Nb, h, Zh)));
    else      WD_TRY((dw_partial<A_GATHER_DIFF, ACT, false>(c, s2, dZ, ld, y, h, 0, Pt, Nb, h, Zh))); }
    // dS = dZ_t * W_h ; dA[a] = sum over out-edges of a of dS  (transpose-CSR, no atomics)
    AOperand a_z{dZ, nullptr, nullptr, nullptr, ld, 0, 0};
    { ProfScope ps(WD_PROF_DX);
    WD_TRY((gemm<A_PLAIN, B_KN, E_STORE, ACT, false>(c, s, a_z, w.W_h, h, nullptr, nullptr, 0, dS, ld, Nb, h, h, &L.pWhT))); }
    ProfScope ps(WD_PROF_SCATTER);
    WD_TRY((aggregate<ACT, false>(s, g, false, dS, ld, nullptr, dA)));
  }
  if (d > 1) {
    { ProfScope ps(WD_PROF_DW);
    WD_TRY(red.add(Ph, gr.dW_h, (int64_t)h * h, (d - 1) * Zh));
    if (c.bias && gr.db_h) WD_TRY(colsum(s, sum, ld, Pcol + pcol, gr.db_h, Nb, h, red)); }  // sum == sum_{t>=1} dZ_t here
    WD_TRY((message_grad<ACT, true, true, false, true>(c, s, g, dA, dS, I, ld, nullptr, sum, dI, Dropout{}, sign_bits(L, ws, Nb, 0, L.pWi), sum2)));
  } else {
    WD_CUDA_OK(cudaMemsetAsync(gr.dW_h, 0, sizeof(float) * h * h, s));
    if (c.bias && gr.db_h) WD_CUDA_OK(cudaMemsetAsync(gr.db_h, 0, sizeof(float) * h, s));
    WD_TRY((message_grad<ACT, true, false, false, true>(c, s, g, dA, dS, I, ld, nullptr, nullptr, dI, Dropout{}, sign_bits(L, ws, Nb, 0, L.pWi))));
  }
  {  // dW_i = dI^T f_bonds ; db_i = colsum(dI)
    ProfScope ps(WD_PROF_DW);
    const int Z = L.Zi;
    AOperand y{g.f_bonds, nullptr, nullptr, nullptr, g.ld_bonds ? g.ld_bonds : Fb, 0, 0};
    WD_TRY((dw_partial<A_PLAIN, ACT, false>(c, s, dI, ld, y, Fb, 0, Pi, Nb, h, Z)));
    WD_TRY(red.add(Pi, gr.dW_i, (int64_t)h * Fb, Z));
    if (c.bias && gr.db_i) WD_TRY(colsum(s, dI, ld, Pcol + 2 * pcol, gr.db_i, Nb, h, red));
  }
  WD_TRY(sd.join());                   // every partial is written before the one reduction launch
  return red.run(s);
}

// ---- small batches: one CUDA graph launch per call ---------------------------------------------------------------
// At the reference's default batch size (50 graphs, chemprop/args.py:95) a forward + backward pass is ~14 us of HBM work
// behind 26 kernel launches: launch gaps, not the kernels, are the step.  For small batches the call's launches are
// therefore CAPTURED (stream capture: nothing runs) into a graph, an executable graph of the same topology is re-used
// through cudaGraphExecUpdate -- pointers, sizes, grids and tensor maps of the new batch replace the old kernel
// parameters, no re-instantiation -- and the whole pass goes to the GPU as ONE launch.  The cache (one executable
// graph per device x pass x configuration) is the only state the library keeps besides the error string; it never
// owns a caller buffer.  Skipped when the caller's stream is itself being captured, under wd_profile_enable (per-launch
// events), for large batches (nothing to gain: WD_GRAPH_MAX_BONDS, default 65536) and with WD_FLAG_NO_GRAPH.
constexpr int kRetryEagerly = 1;      // internal: GraphRun::finish could not capture; nothing has run yet
struct GraphRun {
  cudaStream_t s;                  // the caller's stream: the graph is LAUNCHED here
  cudaStream_t cs = nullptr;       // the stream the pass is captured on: a private one, because the caller's stream may be
  bool active = false;             // the legacy default stream (torch's default), which cannot be captured
  cudaStream_t stream() const { return active ? cs : s; }      // what the pass must launch on
  static cudaStream_t capture_stream() {                       // one per host thread and device, created on first use
    thread_local cudaStream_t st[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return nullptr; }
    if (st[dev] == nullptr && cudaStreamCreateWithFlags(&st[dev], cudaStreamNonBlocking) != cudaSuccess) {
      cudaGetLastError();
      st[dev] = nullptr;
    }
    return st[dev];
  }
  // the side branch of a captured pass (SideBranch above): a second private stream and the events that tie it in
  SideBranch side_branch() const {
    SideBranch sb;
    if (!active) return sb;
    thread_local cudaStream_t st[64] = {};
    thread_local cudaEvent_t ev[64][8] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return sb; }
    if (st[dev] == nullptr) {
      if (cudaStreamCreateWithFlags(&st[dev], cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); st[dev] = nullptr; return sb; }
      for (int i = 0; i < 8; ++i)
        if (cudaEventCreateWithFlags(&ev[dev][i], cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return sb; }
    }
    for (int i = 0; i < 8; ++i) {
      if (ev[dev][i] == nullptr) return sb;
      sb.ev[i] = ev[dev][i];
    }
    sb.side = st[dev];
    return sb;
  }
  using Key = std::array<int64_t, 14>;
  static std::mutex& mu() { static std::mutex m; return m; }
  static std::map<Key, cudaGraphExec_t>& cache() { static std::map<Key, cudaGraphExec_t> c; return c; }
  static int64_t max_bonds() {
    static int64_t v = [] { const char* e = getenv("WD_GRAPH_MAX_BONDS"); return e ? atoll(e) : 65536ll; }();
    return v;
  }

  GraphRun(cudaStream_t stream, const WdConfig& c, const WdGraph& g) : s(stream) {
    if ((c.flags & WD_FLAG_NO_GRAPH) || g_prof_on || g.n_bonds > max_bonds()) return;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) { cudaGetLastError(); return; }
    cs = capture_stream();
    if (cs == nullptr) return;
    if (cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return; }
    active = true;
  }
  // Ends the capture and launches it.  rc: the result of the captured call (its launches were only recorded).
  int finish(int rc, int kind, const WdConfig& c, const WdGraph& g) {
    if (!active) return rc;
    active = false;
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(cs, &graph);
    if (e != cudaSuccess || graph == nullptr) {     // something in the pass was not capturable: the caller runs it eagerly
      cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      return rc != WD_OK ? rc : kRetryEagerly;
    }
    if (rc != WD_OK) { cudaGraphDestroy(graph); return rc; }
    int dev = 0;
    cudaGetDevice(&dev);
    const Key key = {dev, kind, c.hidden_size, c.atom_fdim, c.bond_fdim, c.depth, c.activation, c.aggregation, c.undirected,
                     c.bias, c.math_mode, (int64_t)c.flags, c.dropout > 0.f ? 1 : 0, g.n_mols > 0 ? 1 : 0};
    std::lock_guard<std::mutex> lock(mu());
    auto& cc = cache();
    cudaGraphExec_t exec = nullptr;
    auto it = cc.find(key);
    if (it != cc.end()) {
      exec = it->second;
      cudaGraphExecUpdateResultInfo info;
      if (cudaGraphExecUpdate(exec, graph, &info) != cudaSuccess) {      // topology changed: build a new executable graph
        cudaGetLastError();
        cudaGraphExecDestroy(exec);
        exec = nullptr;
        cc.erase(it);
      }
    }
    if (exec == nullptr) {
      if (cc.size() >= 512) { for (auto& kv : cc) cudaGraphExecDestroy(kv.second); cc.clear(); }
      const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
      if (ei != cudaSuccess) { cudaGraphDestroy(graph); return fail(WD_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ei)); }
      cc[key] = exec;
      g_graph_builds.fetch_add(1, std::memory_order_relaxed);
    }
    cudaGraphDestroy(graph);
    const cudaError_t el = cudaGraphLaunch(exec, s);
    if (el != cudaSuccess) return fail(WD_ERR_CUDA, "cudaGraphLaunch failed: %s", cudaGetErrorString(el));
    g_graph_launches.fetch_add(1, std::memory_order_relaxed);
    return WD_OK;
  }
  ~GraphRun() {   // an early return between begin and finish must not leave the stream capturing
    if (active) { cudaGraph_t g = nullptr; cudaStreamEndCapture(cs, &g); if (g) cudaGraphDestroy(g); cudaGetLastError(); }
  }
};

}  // namespace
}  // namespace wd

// ---- C ABI --------------------------------------------------------------------------------------
extern "C" size_t wd_workspace_bytes(const WdConfig* cfg, int64_t n_atoms, int64_t n_bonds, int64_t n_mols) {
  if (wd::check_cfg(cfg) != WD_OK || n_atoms < 1 || n_bonds < 1 || n_mols < 0) return 0;
  return (size_t)wd::make_layout(*cfg, n_atoms, n_bonds, n_mols).total * sizeof(float);
}

extern "C" int wd_encoder_forward(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg,
                                  void* workspace, size_t workspace_bytes, float* out, void* stream) {
  using namespace wd;
  WD_TRY(check_cfg(cfg));
  WD_TRY(check_graph(graph, weights, cfg));
  if (!workspace || (graph->n_mols > 0 && !out)) return fail(WD_ERR_INVALID_ARG, "workspace/out is NULL");
  if (workspace_bytes < wd_workspace_bytes(cfg, graph->n_atoms, graph->n_bonds, graph->n_mols))
    return fail(WD_ERR_INVALID_ARG, "workspace too small: %zu < %zu bytes", workspace_bytes,
                wd_workspace_bytes(cfg, graph->n_atoms, graph->n_bonds, graph->n_mols));
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return fail(WD_ERR_INVALID_ARG, "workspace must be 256-byte aligned");
  int rc = WD_OK;
  {
    GraphRun gr((cudaStream_t)stream, *cfg, *graph);
    WD_DISPATCH_ACT(cfg->activation, rc = forward_impl<ACT>(*graph, *weights, *cfg, workspace, out, gr.stream(), false));
    rc = gr.finish(rc, 0, *cfg, *graph);
  }
  if (rc == kRetryEagerly) WD_DISPATCH_ACT(cfg->activation, rc = forward_impl<ACT>(*graph, *weights, *cfg, workspace, out, (cudaStream_t)stream, false));
  return rc;
}

extern "C" size_t wd_infer_workspace_bytes(const WdConfig* cfg, int64_t n_atoms, int64_t n_bonds, int64_t n_mols) {
  if (wd::check_cfg(cfg) != WD_OK || n_atoms < 1 || n_bonds < 1 || n_mols < 0) return 0;
  return (size_t)wd::make_layout(*cfg, n_atoms, n_bonds, n_mols, true).total * sizeof(float);
}

extern "C" int wd_encoder_infer(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg,
                                void* workspace, size_t workspace_bytes, float* out, void* stream) {
  using namespace wd;
  WD_TRY(check_cfg(cfg));
  WD_TRY(check_graph(graph, weights, cfg));
  if (!workspace || (graph->n_mols > 0 && !out)) return fail(WD_ERR_INVALID_ARG, "workspace/out is NULL");
  if (workspace_bytes < wd_infer_workspace_bytes(cfg, graph->n_atoms, graph->n_bonds, graph->n_mols))
    return fail(WD_ERR_INVALID_ARG, "workspace too small: %zu < %zu bytes", workspace_bytes,
                wd_infer_workspace_bytes(cfg, graph->n_atoms, graph->n_bonds, graph->n_mols));
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return fail(WD_ERR_INVALID_ARG, "workspace must be 256-byte aligned");
  int rc = WD_OK;
  {
    GraphRun gr((cudaStream_t)stream, *cfg, *graph);
    WD_DISPATCH_ACT(cfg->activation, rc = forward_impl<ACT>(*graph, *weights, *cfg, workspace, out, gr.stream(), true));
    rc = gr.finish(rc, 1, *cfg, *graph);
  }
  if (rc == kRetryEagerly) WD_DISPATCH_ACT(cfg->activation, rc = forward_impl<ACT>(*graph, *weights, *cfg, workspace, out, (cudaStream_t)stream, true));
  return rc;
}

extern "C" int wd_encoder_backward(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg,
                                   void* workspace, size_t workspace_bytes, const float* d_out,
                                   const WdGrads* grads, void* stream) {
  using namespace wd;
  WD_TRY(check_cfg(cfg));
  WD_TRY(check_graph(graph, weights, cfg));
  if (!workspace || !grads || !grads->dW_i || !grads->dW_h || !grads->dW_o || !grads->db_o ||
      (graph->n_mols > 0 && !d_out))
    return fail(WD_ERR_INVALID_ARG, "workspace/grads/d_out is NULL");
  if (workspace_bytes < wd_workspace_bytes(cfg, graph->n_atoms, graph->n_bonds, graph->n_mols))
    return fail(WD_ERR_INVALID_ARG, "workspace too small");
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return fail(WD_ERR_INVALID_ARG, "workspace must be 256-byte aligned");
  int rc = WD_OK;
  {
    GraphRun gr((cudaStream_t)stream, *cfg, *graph);
    const SideBranch sb = (cfg->flags & WD_FLAG_NO_SIDE_BRANCH) ? SideBranch{} : gr.side_branch();
    WD_DISPATCH_ACT(cfg->activation, rc = backward_impl<ACT>(*graph, *weights, *cfg, workspace, d_out, *grads, gr.stream(), sb));
    rc = gr.finish(rc, 2, *cfg, *graph);
  }
  if (rc == kRetryEagerly) WD_DISPATCH_ACT(cfg->activation, rc = backward_impl<ACT>(*graph, *weights, *cfg, workspace, d_out, *grads, (cudaStream_t)stream));
  return rc;
}

extern "C" const char* wd_last_error(void) { return wd::g_err; }
extern "C" int64_t wd_launch_count(void) { return wd::g_launches.load(); }
extern "C" void wd_graph_stats(int64_t* launches, int64_t* builds) {
  if (launches) *launches = wd::g_graph_launches.load();
  if (builds) *builds = wd::g_graph_builds.load();
}

// ---- test hooks: the raw contractions behind the encoder, on plain operands ----------------------
// C[M, ldc] = epi(A[M, K] * B^T (+ bias) (+ R)), B(n,k) = b_kn ? W[k*ldb+n] : W[n*ldb+k].
// epi: 0 store, 1 ReLU, 2 ReLU(R + .).  `scratch` holds the packed weight image in TF32 mode.
extern "C" int wd_debug_gemm(const float* A, int lda, const float* W, int ldb, int b_kn, const float* bias,
                             const float* R, int ldr, float* C, int ldc, int64_t M, int N, int K, int epi,
                             int math_mode, void* scratch, size_t scratch_bytes, void* stream) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream;
  WdConfig c{};
  c.math_mode = math_mode;
  // TF32X3: the forward kernel's geometry (tiles up to 304 columns, hi + lo weight images); b_kn is a backward form
  const bool bx3 = math_mode == WD_MATH_BF16X3 && !b_kn, x3 = (math_mode == WD_MATH_TF32X3 && !b_kn) || bx3;
  TcPacked pk = tc_packed_geometry(N, K, 0, x3 ? T4_MAX_BN : T2_MAX_BN, x3 && !bx3 ? 2 : 1, bx3 ? 1 : 0);
  if (math_mode != WD_MATH_FP32) {
    if (!scratch || scratch_bytes < (size_t)pk.floats() * sizeof(float))
      return fail(WD_ERR_INVALID_ARG, "wd_debug_gemm: scratch needs %lld bytes", (long long)pk.floats() * 4);
    pk.img = reinterpret_cast<float*>(scratch);
    WD_TRY(tc_pack(s, W, ldb, b_kn ? B_KN : B_NK, pk));
  }
  AOperand a{A, nullptr, nullptr, nullptr, lda, 0, K};
  constexpr int ACT = WD_ACT_RELU;
  if (b_kn) {
    if (epi != 0) return fail(WD_ERR_UNSUPPORTED, "wd_debug_gemm: b_kn supports epi 0 only");
    return gemm<A_PLAIN, B_KN, E_STORE, ACT, false>(c, s, a, W, ldb, bias, R, ldr, C, ldc, M, N, K, &pk);
  }
  switch (epi) {
    case 0: return gemm<A_PLAIN, B_NK, E_STORE, ACT, false>(c, s, a, W, ldb, bias, R, ldr, C, ldc, M, N, K, &pk);
    case 1: return gemm<A_PLAIN, B_NK, E_ACT, ACT, false>(c, s, a, W, ldb, bias, R, ldr, C, ldc, M, N, K, &pk);
    case 2: return gemm<A_PLAIN, B_NK, E_RES_ACT, ACT, false>(c, s, a, W, ldb, bias, R, ldr, C, ldc, M, N, K, &pk);
  }
  return fail(WD_ERR_INVALID_ARG, "wd_debug_gemm: epi %d unknown", epi);
}

// out[N1, N2] = X[R, :N1]^T * Y[R, :N2]  (deterministic split over rows + fixed-order reduction).
// `scratch` holds the row-chunk partials.
extern "C" int wd_debug_dw(const float* X, int ldx, const float* Y, int ldy, float* out, int64_t R, int N1, int N2,
                           int math_mode, void* scratch, size_t scratch_bytes, void* stream) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream;
  WdConfig c{};
  c.math_mode = math_mode;
  const int Z = math_mode != WD_MATH_FP32 ? tc_dw_z(R, N1, tc_dw_count_tiles(N2, 0)) : dw_chunks(R, N1, N2);
  if (!scratch || scratch_bytes < (size_t)Z * N1 * N2 * sizeof(float))
    return fail(WD_ERR_INVALID_ARG, "wd_debug_dw: scratch needs %lld bytes", (long long)Z * N1 * N2 * 4);
  float* P = reinterpret_cast<float*>(scratch);
  AOperand y{Y, nullptr, nullptr, nullptr, ldy, 0, 0};
  WD_TRY((dw_partial<A_PLAIN, WD_ACT_RELU, false>(c, s, X, ldx, y, N2, 0, P, R, N1, Z)));
  return reduce_partials(s, P, out, (int64_t)N1 * N2, Z);
}

// Reads (and optionally clears) the per-role wait-cycle counters of CTA 0 of the tcgen05 GEMM kernel
// (all zero unless the library was built with -DWD_T2_TIMING).
extern "C" int wd_debug_timing(uint64_t* out16, int reset) {
  unsigned long long h[16];
  if (cudaMemcpyFromSymbol(h, wd::g_t2_timing, sizeof(h)) != cudaSuccess) return wd::fail(WD_ERR_CUDA, "wd_debug_timing: copy failed");
  for (int i = 0; i < 16; ++i) out16[i] = h[i];
  if (reset) {
    unsigned long long z[16] = {};
    if (cudaMemcpyToSymbol(wd::g_t2_timing, z, sizeof(z)) != cudaSuccess) return wd::fail(WD_ERR_CUDA, "wd_debug_timing: reset failed");
  }
  return 0;
}
