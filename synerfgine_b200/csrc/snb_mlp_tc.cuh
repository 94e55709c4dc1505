// synerfgine_b200 — the two width-64 MLPs of NerfNetwork on the 5th-generation tensor cores (tcgen05 / UMMA), CTA level.
//
// Replaces (reference): tcnn kernel_mlp_fused<64, ..., __half, ReLU, INFERENCE> x2 (dependencies/tiny-cuda-nn/src/
// fully_fused_mlp.cu:499-557, wmma m16n16k16 with fp16 accumulators) plus the wiring of NerfNetwork::
// inference_mixed_precision_impl / extract_density (include/neural-graphics-primitives/nerf_network.h:31-43,105-139).
//
// One CTA = 128 threads = one 128-sample tile (UMMA M = 128, thread r owns sample row r = TMEM lane r):
//   density: [128x32] x W1^T[32x64] -> ReLU -> x W2^T[64x16]            (tcgen05.mma kind::f16, fp32 accumulators in TMEM)
//   rgb:     [128x(16 density out | 16 SH)] x W3^T -> ReLU -> x W4^T[64x64] -> ReLU -> x W5^T[64x16]
// Operands live in shared memory in the canonical K-major, no-swizzle core-matrix layout (8 rows x 16 bytes contiguous;
// LBO = 128 B between K-adjacent core matrices, SBO = (K/8) * 128 B between 8-row groups). All five weight matrices (20 KB)
// are laid out that way once on the host and arrive with ONE bulk TMA copy (cp.async.bulk ... mbarrier::complete_tx) per CTA.
// Between layers each thread pulls its row out of TMEM (tcgen05.ld 32x32b), applies ReLU, rounds to fp16 and writes the next
// A operand; accumulators never touch registers of other threads or HBM. TMEM budget: 128 columns per CTA
// (D1 @0..63, D2 @64..79, D3 @0..63, D4 @64..127, D5 @0..15) so four CTAs share an SM's 512 columns.
//
// Numerics: fp16 operands, fp32 accumulation, every layer output rounded to fp16 (ReLU before rounding is the same as
// after) — identical to the warp-level mma.sync evaluator up to fp32 summation order.
#pragma once
#include "snb_network.cuh"

namespace snb {
namespace tc {

constexpr int kTileRows = 128;
constexpr uint32_t kTmemCols = 128;
// shared-memory carve-up (bytes), 1024-aligned base
constexpr int kSmemW = 0;                                  // 20480: W1 W2 W3 W4 W5 in core-matrix order (offsets = kOff* halves)
constexpr int kSmemA32 = kSmemW + kMlpHalves * 2;          //  8192: [128 x 32] features (density input)
constexpr int kSmemA64 = kSmemA32 + kTileRows * 32 * 2;    // 16384: [128 x 64] hidden activations
constexpr int kSmemR32 = kSmemA64 + kTileRows * 64 * 2;    //  8192: [128 x 32] rgb input (16 density outputs | 16 SH)
constexpr int kSmemBar = kSmemR32 + kTileRows * 32 * 2;    //    32: two mbarriers + TMEM base address
constexpr int kSmemBytes = kSmemBar + 32;

// ---- host: weights -> K-major core-matrix order --------------------------------------------------------------------------
// element (n, k) of a [N x K] row-major matrix goes to half offset (n/8) * (K/8) * 64 + (k/8) * 64 + (n%8) * 8 + (k%8).
inline void build_umma_weights(const uint16_t* mlp_rowmajor, uint16_t* out) {
	struct L { int off, N, K; };
	const L layers[5] = {{kOffDW1, 64, 32}, {kOffDW2, 16, 64}, {kOffRW1, 64, 32}, {kOffRW2, 64, 64}, {kOffRW3, 16, 64}};
	for (const L& l : layers)
		for (int n = 0; n < l.N; ++n)
			for (int k = 0; k < l.K; ++k)
				out[l.off + (n / 8) * (l.K / 8) * 64 + (k / 8) * 64 + (n % 8) * 8 + (k % 8)] = mlp_rowmajor[l.off + n * l.K + k];
}

// ---- PTX wrappers -----------------------------------------------------------------------------------------------------------
SNB_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Shared-memory matrix descriptor (sm_100): [13:0] start >> 4, [29:16] LBO >> 4, [45:32] SBO >> 4, [47:46] version = 1,
// [63:61] layout type 0 = no swizzle.
SNB_D uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
	return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | ((uint64_t)1 << 46);
}
// Instruction descriptor, kind::f16: [5:4] D format 1 = f32, [9:7] A format 0 = f16, [12:10] B format 0 = f16,
// [15] A major 0 = K, [16] B major 0 = K, [22:17] N >> 3, [28:24] M >> 4.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) { return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24); }

SNB_D void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
	             :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
SNB_D void mma_commit(uint32_t mbar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" :: "r"(mbar) : "memory"); }
SNB_D void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
SNB_D void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
SNB_D void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
SNB_D void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
SNB_D void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_dst), "r"(ncols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
SNB_D void tmem_dealloc(uint32_t taddr, uint32_t ncols) { asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(taddr), "r"(ncols) : "memory"); }
SNB_D void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
	               "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr) : "memory");
}
SNB_D void tmem_ld4(uint32_t taddr, uint32_t (&r)[4]) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
SNB_D void mbar_init(uint32_t mbar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(mbar), "r"(count) : "memory"); }
SNB_D void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
SNB_D void mbar_expect_tx(uint32_t mbar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" :: "r"(mbar), "r"(bytes) : "memory"); }
// Bounded spin: a lost arrival (wrong descriptor, faulted MMA) must trap instead of hanging the GPU.
SNB_D void mbar_wait(uint32_t mbar, uint32_t parity) {
	uint32_t done = 0;
	for (uint32_t spins = 0; !done; ++spins) {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
		if (spins > (1u << 24)) __trap();
	}
}
// 1-D bulk TMA copy global -> shared, completion on an mbarrier (SASS: UBLKCP)
SNB_D void tma_bulk_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t mbar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" :: "r"(smem_dst), "l"(gsrc), "r"(bytes), "r"(mbar) : "memory");
}

// byte offset of the 16-byte chunk (row r, K-chunk kc) in a [128 x K] K-major core-matrix tile
template <int K> SNB_D int chunk_offset(int r, int kc) { return (r >> 3) * (K / 8) * 128 + kc * 128 + (r & 7) * 16; }

// D[tmem] = A[128 x K] (smem) * W^T (smem, [N x K]); one elected thread issues K/16 MMAs and commits them to `mbar`.
template <int K, int N>
SNB_D void issue_layer(uint32_t d_tmem, uint32_t a_smem, uint32_t w_smem, uint32_t mbar) {
	constexpr uint32_t sbo = (K / 8) * 128;
#pragma unroll
	for (int s = 0; s < K / 16; ++s)
		mma_f16_ss(d_tmem, make_smem_desc(a_smem + s * 256, 128, sbo), make_smem_desc(w_smem + s * 256, 128, sbo), make_idesc(kTileRows, N), s > 0 ? 1u : 0u);
	mma_commit(mbar);
}

// Epilogue of a 64-wide hidden layer: this thread's row, 4 x 16 fp32 columns -> ReLU -> fp16 -> A operand [128 x 64].
SNB_D void hidden_to_smem(uint32_t taddr, unsigned char* a64, int row) {
#pragma unroll
	for (int c = 0; c < 4; ++c) {
		uint32_t v[16];
		tmem_ld16(taddr + c * 16, v);
		tmem_wait_ld();
		uint32_t p[8];
#pragma unroll
		for (int i = 0; i < 8; ++i) p[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), true);
		*reinterpret_cast<uint4*>(a64 + chunk_offset<64>(row, 2 * c)) = make_uint4(p[0], p[1], p[2], p[3]);
		*reinterpret_cast<uint4*>(a64 + chunk_offset<64>(row, 2 * c + 1)) = make_uint4(p[4], p[5], p[6], p[7]);
	}
}

// Shared state of one CTA's evaluator.
struct CtaMlp {
	unsigned char* smem;       // 1024-aligned base of the kSmemBytes carve-up
	uint32_t tmem;             // TMEM base (lane 0, first allocated column)
	uint32_t bar_w, bar_mma;   // mbarrier addresses: weights landed / MMAs of the current layer done
	uint32_t phase;            // parity of bar_mma
};

// One-time set-up, called by all 128 threads: mbarriers, TMEM allocation, weight TMA. Returns with the weights resident.
SNB_D void cta_mlp_init(CtaMlp& m, unsigned char* smem, const __half* __restrict__ w_umma_global) {
	m.smem = smem;
	m.bar_w = smem_u32(smem + kSmemBar);
	m.bar_mma = smem_u32(smem + kSmemBar + 8);
	m.phase = 0;
	const int tid = threadIdx.x;
	if (tid == 0) {
		mbar_init(m.bar_w, 1); mbar_init(m.bar_mma, 1);
		mbar_fence_init();
	}
	if (tid < 32) tmem_alloc(smem_u32(smem + kSmemBar + 16), kTmemCols);
	fence_before_sync();
	__syncthreads();
	fence_after_sync();
	m.tmem = *reinterpret_cast<volatile uint32_t*>(smem + kSmemBar + 16);
	if (tid == 0) {
		mbar_expect_tx(m.bar_w, kMlpHalves * 2);
		tma_bulk_g2s(smem_u32(smem + kSmemW), w_umma_global, kMlpHalves * 2, m.bar_w);
	}
	mbar_wait(m.bar_w, 0);
}
SNB_D void cta_mlp_free(CtaMlp& m) {
	fence_before_sync();
	__syncthreads();
	if (threadIdx.x < 32) tmem_dealloc(m.tmem, kTmemCols);
}

// hand the A operand just written (generic proxy) to the tensor core (async proxy); thread 0 then issues the layer
SNB_D void publish_operand() { fence_before_sync(); fence_async_smem(); __syncthreads(); }
SNB_D void await_layer(CtaMlp& m) { mbar_wait(m.bar_mma, m.phase); m.phase ^= 1u; fence_after_sync(); }

// Both MLPs for the CTA's 128 rows. Preconditions: every thread wrote its 32 feature halves into the A32 tile and its 16 SH
// halves into chunks 2,3 of the R32 tile (chunk_offset<32>(row, kc)). Returns this thread's {r, g, b, sigma} as raw fp16
// network outputs (same contract as warp_network_eval). All 128 threads must call it.
SNB_D uint2 cta_network_eval(CtaMlp& m) {
	const int row = threadIdx.x;
	const bool issuer = threadIdx.x == 0;
	const uint32_t lane_base = m.tmem + ((uint32_t)(row & ~31) << 16);   // this warp's TMEM lane quarter
	unsigned char* a64 = m.smem + kSmemA64;
	unsigned char* r32 = m.smem + kSmemR32;
	const uint32_t sW = smem_u32(m.smem + kSmemW), sA32 = smem_u32(m.smem + kSmemA32), sA64 = smem_u32(a64), sR32 = smem_u32(r32);

	// density layer 1: 32 -> 64, ReLU                                 D1 @ cols 0..63
	publish_operand();
	if (issuer) { fence_after_sync(); issue_layer<32, 64>(m.tmem + 0, sA32, sW + kOffDW1 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base + 0, a64, row);
	// density layer 2: 64 -> 16, no activation                        D2 @ cols 64..79
	publish_operand();
	if (issuer) { fence_after_sync(); issue_layer<64, 16>(m.tmem + 64, sA64, sW + kOffDW2 * 2, m.bar_mma); }
	await_layer(m);
	uint32_t sigma_bits;
	{
		uint32_t v[16];
		tmem_ld16(lane_base + 64, v);
		tmem_wait_ld();
		uint32_t p[8];
#pragma unroll
		for (int i = 0; i < 8; ++i) p[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), false);
		sigma_bits = p[0] & 0xFFFFu;                                   // density output 0 = log-density (extract_density)
		*reinterpret_cast<uint4*>(r32 + chunk_offset<32>(row, 0)) = make_uint4(p[0], p[1], p[2], p[3]);
		*reinterpret_cast<uint4*>(r32 + chunk_offset<32>(row, 1)) = make_uint4(p[4], p[5], p[6], p[7]);
	}
	// rgb layer 1: [16 density | 16 SH] -> 64, ReLU                   D3 @ cols 0..63
	publish_operand();
	if (issuer) { fence_after_sync(); issue_layer<32, 64>(m.tmem + 0, sR32, sW + kOffRW1 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base + 0, a64, row);
	// rgb layer 2: 64 -> 64, ReLU                                     D4 @ cols 64..127
	publish_operand();
	if (issuer) { fence_after_sync(); issue_layer<64, 64>(m.tmem + 64, sA64, sW + kOffRW2 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base + 64, a64, row);                          // the MMA that read a64 has completed: safe to overwrite
	// rgb layer 3: 64 -> 16 (3 live outputs)                          D5 @ cols 0..15
	publish_operand();
	if (issuer) { fence_after_sync(); issue_layer<64, 16>(m.tmem + 0, sA64, sW + kOffRW3 * 2, m.bar_mma); }
	await_layer(m);
	uint32_t v[4];
	tmem_ld4(lane_base + 0, v);
	tmem_wait_ld();
	uint2 r;
	r.x = h22u(__floats2half2_rn(__uint_as_float(v[0]), __uint_as_float(v[1])));
	r.y = (uint32_t)__half_as_ushort(__float2half_rn(__uint_as_float(v[2]))) | (sigma_bits << 16);
	return r;
}

// ---- compact form for the march kernel: ONE 16 KB operand buffer and 64 TMEM columns per CTA ---------------------------------
// Every layer's output overwrites the operand buffer only after the MMA that read it has completed (await_layer), and every
// accumulator lives at columns 0..63 because layer i+1's MMA is issued only after all threads pulled layer i out of TMEM.
constexpr uint32_t kTmemColsCompact = 64;
constexpr int kCompactW = 0;                                   // 20480 B weights
constexpr int kCompactBuf = kMlpHalves * 2;                    // 16384 B operand buffer ([128 x 32] or [128 x 64])
constexpr int kCompactBar = kCompactBuf + kTileRows * 64 * 2;  // mbarriers + TMEM address
constexpr int kCompactBytes = kCompactBar + 32;

SNB_D void cta_mlp_init_compact(CtaMlp& m, unsigned char* smem, const __half* __restrict__ w_umma_global) {
	m.smem = smem;
	m.bar_w = smem_u32(smem + kCompactBar);
	m.bar_mma = smem_u32(smem + kCompactBar + 8);
	m.phase = 0;
	const int tid = threadIdx.x;
	if (tid == 0) { mbar_init(m.bar_w, 1); mbar_init(m.bar_mma, 1); mbar_fence_init(); }
	if (tid < 32) tmem_alloc(smem_u32(smem + kCompactBar + 16), kTmemColsCompact);
	fence_before_sync();
	__syncthreads();
	fence_after_sync();
	m.tmem = *reinterpret_cast<volatile uint32_t*>(smem + kCompactBar + 16);
	if (tid == 0) {
		mbar_expect_tx(m.bar_w, kMlpHalves * 2);
		tma_bulk_g2s(smem_u32(smem + kCompactW), w_umma_global, kMlpHalves * 2, m.bar_w);
	}
	mbar_wait(m.bar_w, 0);
}
SNB_D void cta_mlp_free_compact(CtaMlp& m) {
	fence_before_sync();
	__syncthreads();
	if (threadIdx.x < 32) tmem_dealloc(m.tmem, kTmemColsCompact);
}

// Precondition: every thread wrote its 32 feature halves into the buffer in the [128 x 32] layout (chunk_offset<32>(row, 0..3)).
// sh = this row's 16 SH halves. Returns {r, g, b, sigma} raw fp16 network outputs. All 128 threads must call it.
SNB_D uint2 cta_network_eval_compact(CtaMlp& m, const uint32_t (&sh)[8]) {
	const int row = threadIdx.x;
	const bool issuer = threadIdx.x == 0;
	const uint32_t lane_base = m.tmem + ((uint32_t)(row & ~31) << 16);
	unsigned char* buf = m.smem + kCompactBuf;
	const uint32_t sW = smem_u32(m.smem + kCompactW), sBuf = smem_u32(buf);
	publish_operand();                                                                  // density 32 -> 64, ReLU
	if (issuer) { fence_after_sync(); issue_layer<32, 64>(m.tmem, sBuf, sW + kOffDW1 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base, buf, row);
	publish_operand();                                                                  // density 64 -> 16
	if (issuer) { fence_after_sync(); issue_layer<64, 16>(m.tmem, sBuf, sW + kOffDW2 * 2, m.bar_mma); }
	await_layer(m);
	uint32_t sigma_bits;
	{
		uint32_t v[16];
		tmem_ld16(lane_base, v);
		tmem_wait_ld();
		uint32_t p[8];
#pragma unroll
		for (int i = 0; i < 8; ++i) p[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), false);
		sigma_bits = p[0] & 0xFFFFu;
		*reinterpret_cast<uint4*>(buf + chunk_offset<32>(row, 0)) = make_uint4(p[0], p[1], p[2], p[3]);
		*reinterpret_cast<uint4*>(buf + chunk_offset<32>(row, 1)) = make_uint4(p[4], p[5], p[6], p[7]);
		*reinterpret_cast<uint4*>(buf + chunk_offset<32>(row, 2)) = make_uint4(sh[0], sh[1], sh[2], sh[3]);
		*reinterpret_cast<uint4*>(buf + chunk_offset<32>(row, 3)) = make_uint4(sh[4], sh[5], sh[6], sh[7]);
	}
	publish_operand();                                                                  // rgb [16 | 16 SH] -> 64, ReLU
	if (issuer) { fence_after_sync(); issue_layer<32, 64>(m.tmem, sBuf, sW + kOffRW1 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base, buf, row);
	publish_operand();                                                                  // rgb 64 -> 64, ReLU
	if (issuer) { fence_after_sync(); issue_layer<64, 64>(m.tmem, sBuf, sW + kOffRW2 * 2, m.bar_mma); }
	await_layer(m);
	hidden_to_smem(lane_base, buf, row);
	publish_operand();                                                                  // rgb 64 -> 16 (3 live)
	if (issuer) { fence_after_sync(); issue_layer<64, 16>(m.tmem, sBuf, sW + kOffRW3 * 2, m.bar_mma); }
	await_layer(m);
	uint32_t v[4];
	tmem_ld4(lane_base, v);
	tmem_wait_ld();
	uint2 r;
	r.x = h22u(__floats2half2_rn(__uint_as_float(v[0]), __uint_as_float(v[1])));
	r.y = (uint32_t)__half_as_ushort(__float2half_rn(__uint_as_float(v[2]))) | (sigma_bits << 16);
	return r;
}

} // namespace tc

// NerfNetwork inference on explicit NerfCoordinates (n x 7 floats) through the tcgen05 evaluator: the parity tap of the
// tensor-core path (same contract as k_network_inference). grid-stride over 128-sample tiles; dynamic smem = tc::kSmemBytes + 2 KB.
template <int F>
__global__ void __launch_bounds__(tc::kTileRows) k_network_inference_tc(DevModel model, const float* __restrict__ coords, uint64_t n, uint16_t* __restrict__ out,
                                                                        uint16_t* __restrict__ features, uint32_t* __restrict__ hash_idx) {
	extern __shared__ unsigned char smem_dyn[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
	LevelDesc* levels = reinterpret_cast<LevelDesc*>(smem + tc::kSmemBytes);
	const int tid = threadIdx.x;
	if (tid < kFeat / F) levels[tid] = make_level_desc(model.grid, tid, model.grid_params);
	tc::CtaMlp mlp;
	tc::cta_mlp_init(mlp, smem, model.mlp_umma);
	__syncthreads();
	const uint64_t n_tiles = (n + tc::kTileRows - 1) / tc::kTileRows;
	for (uint64_t tile_i = blockIdx.x; tile_i < n_tiles; tile_i += gridDim.x) {
		const uint64_t i = tile_i * tc::kTileRows + tid;
		const bool valid = i < n;
		unsigned char* frow = smem + tc::kSmemA32 + tc::chunk_offset<32>(tid, 0);   // chunk kc of this row is at frow + kc * 128
		unsigned char* rrow = smem + tc::kSmemR32 + tc::chunk_offset<32>(tid, 0);
		uint32_t sh[8];
#pragma unroll
		for (int j = 0; j < 8; ++j) sh[j] = 0u;
#pragma unroll
		for (int kc = 0; kc < 4; ++kc) *reinterpret_cast<uint4*>(frow + kc * 128) = make_uint4(0, 0, 0, 0);
		if (valid) {
			const float* c = coords + i * 7;
			const v3 p = mk3(c[0], c[1], c[2]), dir = mk3(c[4], c[5], c[6]);
			if (hash_idx) encode_position_to_row<F, true, 128>(levels, p, reinterpret_cast<__half*>(frow), hash_idx + i * (uint64_t)(kFeat / F) * 8);
			else encode_position_to_row<F, false, 128>(levels, p, reinterpret_cast<__half*>(frow), nullptr);
			encode_direction_sh4(dir, sh);
			if (features) {
#pragma unroll
				for (int kc = 0; kc < 4; ++kc) reinterpret_cast<uint4*>(features + i * 32)[kc] = *reinterpret_cast<const uint4*>(frow + kc * 128);
			}
		}
		*reinterpret_cast<uint4*>(rrow + 2 * 128) = make_uint4(sh[0], sh[1], sh[2], sh[3]);
		*reinterpret_cast<uint4*>(rrow + 3 * 128) = make_uint4(sh[4], sh[5], sh[6], sh[7]);
		const uint2 r = tc::cta_network_eval(mlp);
		if (valid) reinterpret_cast<uint2*>(out)[i] = r;
	}
	tc::cta_mlp_free(mlp);
}

} // namespace snb
