// The element-block linear system object shared by system.cu (construction, operators, BLAS-1, saddle-point PCG),
// schur_cg.cu (Schur-complement CG / MINRES with the fused NVLink peer exchange) and precond.cu.
#pragma once
#include "internal.cuh"

#define SYS_THREADS 256
#define MFV2D_MAX_PEERS 16

// Device view of the peer-to-peer exchange blocks (one cudaMalloc'ed, IPC-exported block per rank; peer.cu).  Two exchange
// TYPES with their own flags, epoch counters and double-buffered data areas: 0 = interface multipliers + a few scalars,
// 1 = the coarse-space residual of the two-level preconditioner.
//   flags[type][MFV2D_MAX_PEERS]                         u64: flags[t][r] = last epoch of type t whose data rank r has written HERE
//   data[type][2 parities][n_ranks senders][stride[t]]   doubles; type 0: contributions by GLOBAL interface slot, then the
//                                                        sender's scalars at [n_iface .. n_iface + 3]; type 1: coarse residual
// base[r] is rank r's block mapped into this process (base[rank] is the local block).
struct PeerView
{
    int n_ranks = 1, rank = 0;
    long long n_iface = 0;
    long long stride[2] = {0, 0};   // doubles per sender region
    long long area_off[2] = {0, 0}; // first double of the type's area, counted from the data start
    unsigned char *base[MFV2D_MAX_PEERS] = {};
    unsigned long long *epoch = nullptr; // device counters [2]: exchanges of each type started by this rank
    int *error = nullptr;                // device flag: set when a wait timed out
    int dbg = 0;                         // MFV2D_PEER_DEBUG bits (timing experiments only): 1 = no per-writer fence,
                                         // 2 = one CTA polls the peers' flags, the others a local flag, 8 = timeline stamps
    unsigned long long *stats = nullptr; // [32] timeline of one rank's stream (dbg & 8): ns accumulated before each stamp id,
                                         // [16] = time of the last stamp, [17] = local "peers arrived" epochs (dbg & 2), two types
};
#define PEER_FLAG_BYTES 512 // flags live in the first 2 * 16 * 8 = 256 bytes, the data area starts here
#define PEER_XSCALARS 4     // scalars that ride behind the interface values

#ifndef MFV2D_EMU
__device__ __forceinline__ void peer_fence() { __threadfence_system(); }
__device__ __forceinline__ unsigned long long peer_load_flag(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// the flag stores follow ONE peer_fence() (fence.sys + relaxed store = release pattern); a st.release.sys per peer would
// serialise one NVLink round trip per rank (measured: +47 us per iteration on 8 GPUs)
__device__ __forceinline__ void peer_store_flag(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double peer_load(const double *p) { return __ldcg(p); } // written by another GPU: never from L1
__device__ __forceinline__ void peer_backoff() { __nanosleep(64); }
#else
void emu_peer_fence();
unsigned long long emu_peer_load_flag(const unsigned long long *p);
void emu_peer_store_flag(unsigned long long *p, unsigned long long v);
void emu_peer_backoff();
static inline void peer_fence() { emu_peer_fence(); }
static inline unsigned long long peer_load_flag(const unsigned long long *p) { return emu_peer_load_flag(p); }
static inline void peer_store_flag(unsigned long long *p, unsigned long long v) { emu_peer_store_flag(p, v); }
static inline double peer_load(const double *p) { return *(const volatile double *)p; }
static inline void peer_backoff() { emu_peer_backoff(); }
#endif

// timeline stamp (ONE thread): the time since the previous stamp of this rank's stream is added to stats[id]
__device__ __forceinline__ void peer_stamp(const PeerView &pv, int id)
{
#ifndef MFV2D_EMU
    if ((pv.dbg & 8) && pv.stats)
    {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        const unsigned long long last = pv.stats[16];
        if (last && now > last)
            pv.stats[id] += now - last;
        pv.stats[16] = now;
        __threadfence();
    }
#endif
}

__device__ __forceinline__ double *peer_region(const PeerView &pv, int holder, int type, int parity, int sender)
{
    return reinterpret_cast<double *>(pv.base[holder] + PEER_FLAG_BYTES) + pv.area_off[type] +
           ((long long)parity * pv.n_ranks + sender) * pv.stride[type];
}
__device__ __forceinline__ unsigned long long *peer_flag(const PeerView &pv, int holder, int type, int sender)
{
    return reinterpret_cast<unsigned long long *>(pv.base[holder]) + type * MFV2D_MAX_PEERS + sender;
}
// wait until every other rank has finished writing epoch `e` of `type` into this rank's block (thread r < n_ranks polls
// rank r's flag); a wait that exceeds the budget sets the error flag instead of hanging the GPU.  Ends with a CTA barrier.
__device__ __forceinline__ void peer_wait_all(const PeerView &pv, int type, unsigned long long e)
{
#ifndef MFV2D_EMU
    if ((pv.dbg & 2) && pv.stats)
    {
        // experiment: only CTA 0 polls the peers' (system-scope) flags; the other CTAs wait for its local release
        unsigned long long *arrived = pv.stats + 17 + type;
        if (blockIdx.x == 0)
        {
            if ((int)threadIdx.x < pv.n_ranks && (int)threadIdx.x != pv.rank && !*(volatile int *)pv.error)
            {
                const unsigned long long *f = peer_flag(pv, pv.rank, type, (int)threadIdx.x);
                unsigned long long spins = 0;
                while (peer_load_flag(f) < e)
                {
                    if (++spins > (1ull << 26))
                    {
                        *pv.error = 1;
                        break;
                    }
                    peer_backoff();
                }
            }
            __syncthreads();
            if (threadIdx.x == 0)
                asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(arrived), "l"(e) : "memory");
        }
        else
        {
            if (threadIdx.x == 0)
            {
                unsigned long long v = 0, spins = 0;
                do
                {
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(arrived) : "memory");
                    if (v >= e || *(volatile int *)pv.error)
                        break;
                    peer_backoff();
                } while (++spins < (1ull << 27));
            }
            __syncthreads();
        }
        return;
    }
#endif
    if ((int)threadIdx.x < pv.n_ranks && (int)threadIdx.x != pv.rank && !*(volatile int *)pv.error)
    {
        const unsigned long long *f = peer_flag(pv, pv.rank, type, (int)threadIdx.x);
        unsigned long long spins = 0;
        while (peer_load_flag(f) < e)
        {
            if (++spins > (1ull << 26))
            {
                *pv.error = 1;
                break;
            }
            peer_backoff();
        }
    }
#ifdef MFV2D_EMU
    // one sequential "thread" per CTA: poll every peer in turn
    for (int r = 0; r < pv.n_ranks; ++r)
        if (r != pv.rank)
            while (peer_load_flag(peer_flag(pv, pv.rank, type, r)) < e)
                peer_backoff();
#endif
    __syncthreads();
}

// per-element descriptor of the packed Schur product: one 16-byte load gives the block, the first slot and the block size
struct __align__(16) SchurElemDesc
{
    long long s_off;
    int s0, m;
};

struct mfv2d_system
{
    mfv2d_ctx *ctx;
    int64_t n_elem = 0, n_dense = 0, n_trace = 0, n_con = 0, n_slots = 0;
    int max_n = 0, max_m = 0;
    bool uniform = false; // all blocks have the same size
    bool scaled = false;    // the condensed blocks are those of D S D, D = d_rs (row equilibration, system.cu)
    double *d_rs = nullptr; // [n_trace] power-of-two row scales (only when scaled)
    bool symmetric = false; // every A_e is symmetric to rounding: S_e is stored packed (half the bytes per application)
    // device arrays
    int64_t *d_block_off = nullptr; // [n_elem+1] dense offsets
    int64_t *d_mat_off = nullptr;   // [n_elem+1] offsets into A / Ainv
    double *d_A = nullptr, *d_Ainv = nullptr;
    bool owns_A = true;
    int64_t *d_n_ptr = nullptr, *d_n_col = nullptr; // N   CSR  [n_trace+1], [n_con]
    double *d_n_val = nullptr;
    int64_t *d_t_ptr = nullptr, *d_t_col = nullptr; // N^T CSR  [n_dense+1], [n_con]
    double *d_t_val = nullptr;
    // per element local constraint rows
    int64_t *d_slot_off = nullptr;   // [n_elem+1] first slot of element e (m_e = difference)
    int64_t *d_slot_trace = nullptr; // [n_slots] multiplier of each slot
    int64_t *d_lrow_ptr = nullptr;   // [n_slots+1] entries of each local row
    int32_t *d_lrow_col = nullptr;   // local column index
    double *d_lrow_val = nullptr;
    SchurElemDesc *d_edesc = nullptr; // [n_elem] (s_off, first slot, m) when the counts fit 32 bits
    int32_t *d_slot_trace32 = nullptr; // [n_slots] 32-bit copy of d_slot_trace
    int64_t *d_s_off = nullptr; // [n_elem+1] offsets into S
    double *d_S = nullptr;      // condensed blocks: column-major m x m, or packed "cyclic half" m (m + 1) / 2 when symmetric
    int64_t *d_tr_ptr = nullptr, *d_tr_slot = nullptr; // per multiplier: contributing slots
    int32_t *d_tr2 = nullptr;                          // [n_trace][2] the same as a fixed-width table (-1: none) when no multiplier has more than two slots
    double *d_slot_val = nullptr;                      // [n_slots] scratch: local results
    double *d_diag_inv = nullptr;                      // [n_trace] 1 / diag(S) (Jacobi preconditioner on the trace space)
    // multilevel preconditioner on the trace space (precond.cu): Jacobi + one vector per aggregate on every level
    int n_clevels = 0;
    long long n_agg_total = 0, exact_begin = 0, n_exact = 0, n_chunks = 0;
    int32_t *d_lagg = nullptr;          // [n_clevels][n_trace] aggregate of each multiplier on each level (-1: none)
    long long *d_agg_off = nullptr;     // [n_clevels] first aggregate of the level in the concatenated numbering
    double *d_cz = nullptr;             // [n_trace] near-null-space vector
    int64_t *d_chunk_ptr = nullptr;     // [n_chunks+1] multipliers of each chunk (<= COARSE_CHUNK of one aggregate) ...
    int64_t *d_chunk_idx = nullptr;     // ... in increasing order
    int64_t *d_aggchunk_ptr = nullptr;  // [n_agg_total+1] chunks of each aggregate
    double *d_cpart = nullptr;          // [n_chunks] partial restrictions
    double *d_ldinv = nullptr;          // [exact_begin] 1 / diag(P_l^T S P_l)
    double *d_cy = nullptr;             // [n_agg_total] scaled restrictions | this rank's part of the exact level's residual
    double *d_yc = nullptr;             // [n_exact] correction of the exact level
    double *d_crs = nullptr;            // [n_exact] restricted residual summed over the ranks (peer exchange)
    double *d_Ac_inv = nullptr;         // [n_exact^2] inverse of P_L^T S P_L, row-major
    bool coarse_ready = false;
    // work vectors
    double *d_work_dense[3] = {nullptr, nullptr, nullptr};
    double *d_work_trace[8] = {};
    double *d_partials = nullptr; // reduction partials
    int n_partials = 0;
    // multi-GPU interface (comm.cu)
    int64_t n_shared = 0, n_iface = 0;
    bool interface_set = false;
    int64_t *d_shared_index = nullptr, *d_shared_slot = nullptr;
    double *d_iface = nullptr; // [n_iface + 8]
    double *d_owned = nullptr; // [n_trace] 1.0 / 0.0 ownership weights for dot products (nullptr: all owned)
    // fused peer exchange (schur_cg.cu): which ranks hold each shared multiplier, mapped peer buffers
    bool peers_ready = false;
    PeerView peer;
    int32_t *d_shared_of = nullptr;  // [n_trace] position in the shared list or -1
    int64_t n_coarse_cap = 0;         // capacity of the type-1 exchange area (doubles)
    int64_t *d_holder_ptr = nullptr; // [n_shared+1]
    int32_t *d_holder_rank = nullptr; // the OTHER ranks holding the shared multiplier, ascending
    void *peer_block = nullptr;       // this rank's exchange block (cudaMalloc, IPC exported)
    std::vector<void *> peer_mapped;  // blocks of the other ranks opened with cudaIpcOpenMemHandle
    // fused-CG device state: [0..31] scalars, counters follow (see schur_cg.cu)
    double *d_cg = nullptr;
    // cached CUDA graph of 16 fused CG iterations (pointers inside are the system's own buffers)
    void *cg_graph_exec = nullptr;
    int cg_graph_key = -1;
    uint64_t cg_graph_launches = 0;
    std::vector<void *> allocs;
};

template <typename T> static int sys_upload(mfv2d_system *s, const std::vector<T> &h, T **d)
{
    void *p = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(s->ctx, (h.size() ? h.size() : 1) * sizeof(T), &p));
    s->allocs.push_back(p);
    *d = (T *)p;
    if (!h.empty())
        MFV2D_TRY(mfv2d_cuda_upload(s->ctx, p, h.data(), h.size() * sizeof(T)));
    return MFV2D_SUCCESS;
}

template <typename T> static int sys_alloc(mfv2d_system *s, size_t n, T **d)
{
    void *p = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(s->ctx, (n ? n : 1) * sizeof(T), &p));
    s->allocs.push_back(p);
    *d = (T *)p;
    return MFV2D_SUCCESS;
}

static inline unsigned grid_for(const mfv2d_ctx *ctx, long long n, int threads)
{
    long long g = (n + threads - 1) / threads;
    const long long cap = (long long)ctx->sm_count * 8;
    if (g > cap)
        g = cap;
    if (g < 1)
        g = 1;
    return (unsigned)g;
}

// packed storage of a symmetric m x m block ("cyclic half"): entry (k, r) at k * m + r holds S[r][(r + k) % m] for
// k = 0 .. m / 2 (the last diagonal of an even m only for r < m / 2): m (m + 1) / 2 doubles, every unordered pair once,
// and for a fixed k the m entries are contiguous (coalesced across lanes r) and hit m distinct rows AND columns
static inline long long packed_size(long long m) { return m * (m + 1) / 2; }

int sys_gemv(const mfv2d_system *s, const double *M, const double *x, double *y);
int sys_spmv(const mfv2d_system *s, long long rows, const int64_t *ptr, const int64_t *col, const double *val, const double *x,
             double *y);
int sys_interface_sum(const mfv2d_system *s, double *t, int extra_scalars, double *scalars);
int sys_apply_trace(const mfv2d_system *s, const double *x, double *t);
int sys_schur_local(const mfv2d_system *s, const double *t_in, const int *done);
int sys_apply_schur(const mfv2d_system *s, const double *t_in, double *t_out, const int *done, bool interface_sum = true);
int sys_schur_diagonal(const mfv2d_system *s, double *out);
int sys_dot_device(const mfv2d_system *s, long long n, const double *x, const double *y, const double *w, double *d_result);
int sys_require_interface(const mfv2d_system *s);
int sys_scale_rows(const mfv2d_system *s, double *v, bool inverse);
bool sys_use_peers(const mfv2d_system *s);
int read_scalars(mfv2d_ctx *ctx, const double *d, int n, double *h);
void sys_release_peers(mfv2d_system *s);
int comm_allreduce(mfv2d_ctx *ctx, double *dptr, size_t n);
// peer exchange (peer.cu): all-reduce of the shared multipliers of t (and n_extra scalars) over mapped peer memory
int peer_interface_sum(const mfv2d_system *s, double *t, int n_extra, double *extra);
int peer_dense_sum(const mfv2d_system *s, double *v, long long n);

__global__ void axpby_kernel(long long n, double a, const double *__restrict__ x, double b, const double *__restrict__ y,
                             double *__restrict__ out);
__global__ void trace_gather_kernel(long long n_trace, const int64_t *__restrict__ tr_ptr, const int64_t *__restrict__ tr_slot,
                                    const double *__restrict__ slot_val, double *__restrict__ out, const int *__restrict__ done);
__global__ void dot_partial_kernel(long long n, const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ w,
                                   double *__restrict__ partials, const int *__restrict__ done);
__global__ void sum_partials_kernel(int n_partials, const double *__restrict__ partials, double *__restrict__ result);

// deterministic two-stage reductions shared by the BLAS-1 and CG kernels
__device__ __forceinline__ void block_reduce_store(double acc, double *partials)
{
    __shared__ double red[32];
    acc = warp_sum(acc);
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    if ((threadIdx.x & 31) == 0)
        red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double s = 0.0;
        for (int w = 0; w < nwarps; ++w)
            s += red[w];
        partials[blockIdx.x] = s;
    }
    __syncthreads();
}

// deterministic sum of per-CTA partials by ONE warp: lane-strided, then a butterfly
__device__ __forceinline__ double warp_sum_partials(int n_partials, const double *__restrict__ partials)
{
    double s = 0.0;
    for (int i = MFV2D_LANE0; i < n_partials; i += MFV2D_LANE_STEP)
        s += partials[i];
    return warp_sum(s);
}
