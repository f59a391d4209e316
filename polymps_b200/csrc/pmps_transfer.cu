// pmps_transfer.cu — pmps_upload / pmps_download / pmps_upload_device: the reference's host layout
// (include/MpsParticle.h:139-241: scalars a[i], vectors a[3*i+k], int32 ids, 1-byte bool) <-> the device layout
// (slot order, 32-byte pos/vel records, one plane per component).  Host arrays are pinned in place on first use
// (cudaHostRegister) so the copies run at PCIe rate without an extra host-side staging copy.
#include "pmps_device.cuh"
#include <cstring>

// `rows` (optional): the n entries of the host-side arrays are the particles with these host indices (a compact row set); without it
// entry t is host index t
__global__ void k_up_f64(const double* __restrict__ src, double* dst, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[slotOfHost[rows ? rows[t] : t]] = src[t];
}
__global__ void k_up_i32(const int32_t* __restrict__ src, int32_t* dst, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[slotOfHost[rows ? rows[t] : t]] = src[t];
}
__global__ void k_up_u8(const uint8_t* __restrict__ src, uint8_t* dst, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[slotOfHost[rows ? rows[t] : t]] = src[t];
}
__global__ void k_up_v3(const double* __restrict__ src, V3 dst, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	const int s = slotOfHost[rows ? rows[t] : t];
	dst.x[s] = src[3 * (size_t)t]; dst.y[s] = src[3 * (size_t)t + 1]; dst.z[s] = src[3 * (size_t)t + 2];
}
// pos/vel records: xyz from an interleaved array, pos.w from particleType / PTYPE (any of the sources may be NULL)
__global__ void k_up_pack(const double* __restrict__ xyz, const int32_t* __restrict__ type, const int32_t* __restrict__ ptype,
                          double4* dst, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	const int s = slotOfHost[rows ? rows[t] : t];
	double4 r = dst[s];
	if (xyz) { r.x = xyz[3 * (size_t)t]; r.y = xyz[3 * (size_t)t + 1]; r.z = xyz[3 * (size_t)t + 2]; }
	if (type || ptype) r.w = pk_make(type ? type[t] : pk_type(r), ptype ? ptype[t] : pk_ptype(r));
	dst[s] = r;
}

// positions replaced from the host while a bucket order and its lists exist: the same bookkeeping as a position update inside a
// step (displacement from the build / hit-list positions), plus a flag when a type word changes — then the caller rebuilds
__global__ void k_up_pos_soft(const double* __restrict__ xyz, const int32_t* __restrict__ type, const int32_t* __restrict__ ptype,
                              const __grid_constant__ DevView S, const __grid_constant__ KParams K, int n0) {
	const int hI = blockIdx.x * blockDim.x + threadIdx.x; if (hI >= n0) return;
	const int s = S.slotOfHost[hI];
	const double4 old = S.pos[s]; double4 r = old;
	r.x = xyz[3 * (size_t)hI]; r.y = xyz[3 * (size_t)hI + 1]; r.z = xyz[3 * (size_t)hI + 2];
	if (type || ptype) r.w = pk_make(type ? type[hI] : pk_type(r), ptype ? ptype[hI] : pk_ptype(r));
	if (__double_as_longlong(r.w) != __double_as_longlong(old.w)) S.flags[FLAG_UPLOAD_TYPES] = 1;
	S.pos[s] = r;
	if (pk_type(r) != K.p.ghost) {
		if (owned_slot(S, s)) track_disp(S, K, s, r);
		else track_hit_disp(S, K, s, r);   // a halo copy (slab rank): its position as the caller has it — does the hit list still cover it?
	}
}

__global__ void k_dn_f64(double* dst, const double* __restrict__ src, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[t] = src[slotOfHost[rows ? rows[t] : t]];
}
__global__ void k_dn_i32(int32_t* dst, const int32_t* __restrict__ src, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[t] = src[slotOfHost[rows ? rows[t] : t]];
}
__global__ void k_dn_u8(uint8_t* dst, const uint8_t* __restrict__ src, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	dst[t] = src[slotOfHost[rows ? rows[t] : t]];
}
__global__ void k_dn_v3(double* dst, V3 src, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	const int s = slotOfHost[rows ? rows[t] : t];
	dst[3 * (size_t)t] = src.x[s]; dst[3 * (size_t)t + 1] = src.y[s]; dst[3 * (size_t)t + 2] = src.z[s];
}
__global__ void k_dn_pack(double* xyz, int32_t* type, int32_t* ptype, const double4* __restrict__ src, const int32_t* __restrict__ slotOfHost, int n, const int32_t* __restrict__ rows) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= n) return;
	const double4 r = src[slotOfHost[rows ? rows[t] : t]];
	if (xyz) { xyz[3 * (size_t)t] = r.x; xyz[3 * (size_t)t + 1] = r.y; xyz[3 * (size_t)t + 2] = r.z; }
	if (type) type[t] = pk_type(r);
	if (ptype) ptype[t] = pk_ptype(r);
}
__global__ void k_set_live(const __grid_constant__ DevView S, int live) { S.flags[FLAG_LIVE] = live; }

// Caller memory is never page-locked behind the caller's back: a transfer from pageable memory is correct (the driver stages it),
// a caller that wants the asynchronous DMA path registers its arrays once (pmps_register_host) and owns their lifetime.
extern "C" int pmps_register_host(pmps_t* h, void* p, size_t bytes) {
	if (!h || !p || !bytes) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	for (auto& r : h->registered) if (r.first == p) {
		if (r.second >= bytes) return PMPS_OK;
		return pmps_fail(h, PMPS_E_ARG, "pmps_register_host: %p is registered with %zu bytes, asked for %zu: unregister it first", p, r.second, bytes);
	}
	const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
	if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return PMPS_OK; }   // pinned by the caller itself: nothing to own
	if (e != cudaSuccess) { cudaGetLastError(); return pmps_fail(h, PMPS_E_CUDA, "pmps_register_host(%p, %zu): %s", p, bytes, cudaGetErrorString(e)); }
	h->registered.push_back({ p, bytes });
	return PMPS_OK;
}
extern "C" int pmps_unregister_host(pmps_t* h, void* p) {
	if (!h || !p) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	for (size_t k = 0; k < h->registered.size(); k++) if (h->registered[k].first == p) {
		cudaStreamSynchronize(h->stream); if (h->ioStream) cudaStreamSynchronize(h->ioStream);   // no copy may still be using the pages
		cudaHostUnregister(p); h->registered.erase(h->registered.begin() + k); return PMPS_OK;
	}
	return PMPS_OK;   // not ours (never registered here, or pinned by the caller)
}

static int upload_impl(pmps_handle* h, const PmpsHostArrays* a, bool deviceSrc, bool sameRows = false) {
	if (!h || !a) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	{ const int rc = pmps_ensure_capacity(h, a->numParticlesZero); if (rc) return rc; }   // first upload fixes the capacity
	if (a->numParticles < 0 || a->numParticles > a->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_upload: bad numParticles");
	const int n0 = h->numParticlesZero, T = 256, G = cdiv(n0, T);
	DevView& S = h->V; char* st = (char*)h->stage; const int32_t* soh = S.slotOfHost;
	// stage one host array (or pass a device pointer through)
	auto src = [&](const void* p, size_t bytes, size_t off) -> const void* {
		if (!p) return nullptr;
		if (deviceSrc) return p;
		cudaMemcpyAsync(st + off, p, bytes, cudaMemcpyHostToDevice, h->stream);
		return st + off;
	};
	bool soft = false;
	if (a->pos || a->particleType || a->PTYPE) {
		const double* x = (const double*)src(a->pos, 24 * (size_t)n0, 0);
		const int32_t* t = (const int32_t*)src(a->particleType, 4 * (size_t)n0, 24 * (size_t)n0);
		const int32_t* pt = (const int32_t*)src(a->PTYPE, 4 * (size_t)n0, 28 * (size_t)n0);
		// (a slab rank's table changes its rows at a re-partition: only the caller knows when it hands back the SAME rows, pmps_upload_values_device)
		soft = a->pos && h->haveList && (!h->slab || sameRows) && a->numParticles == h->numParticles && !getenv("PMPS_HARD_UPLOAD");
		if (soft) { cudaMemsetAsync(&S.flags[FLAG_UPLOAD_TYPES], 0, sizeof(int32_t), h->stream); LAUNCH(h, "upload", k_up_pos_soft, G, T, x, t, pt, S, h->K, n0); }
		else LAUNCH(h, "upload", k_up_pack, G, T, x, t, pt, S.pos, soh, n0, (const int32_t*)nullptr);
	}
	if (a->vel) { const double* x = (const double*)src(a->vel, 24 * (size_t)n0, 0); LAUNCH(h, "upload", k_up_pack, G, T, x, (const int32_t*)nullptr, (const int32_t*)nullptr, S.vel, soh, n0, (const int32_t*)nullptr); }
	if (a->particleNearWall) { const uint8_t* x = (const uint8_t*)src(a->particleNearWall, (size_t)n0, 0); LAUNCH(h, "upload", k_up_u8, G, T, x, S.particleNearWall, soh, n0, (const int32_t*)nullptr); }
#define X(n) if (a->n) { const int32_t* x = (const int32_t*)src(a->n, 4 * (size_t)n0, 0); LAUNCH(h, "upload", k_up_i32, G, T, x, S.n, soh, n0, (const int32_t*)nullptr); }
	PMPS_I32_FIELDS(X)
#undef X
#define X(n) if (a->n) { const double* x = (const double*)src(a->n, 8 * (size_t)n0, 0); LAUNCH(h, "upload", k_up_f64, G, T, x, S.n, soh, n0, (const int32_t*)nullptr); }
	PMPS_F64_FIELDS(X)
	if (h->K.p.fluidType == 1) { PMPS_RHEO_FIELDS(X) }
#undef X
#define X(n) if (a->n) { const double* x = (const double*)src(a->n, 24 * (size_t)n0, 0); LAUNCH(h, "upload", k_up_v3, G, T, x, S.n, soh, n0, (const int32_t*)nullptr); }
	PMPS_V3_FIELDS(X)
#undef X
	LAUNCH(h, "upload", k_set_live, 1, 1, S, a->numParticles);
	CUDA_TRY(h, cudaGetLastError());
	h->numParticles = a->numParticles; h->haveState = true;
	if (a->pos) h->guardWorst = 0.0;   // a new state from the caller: the sticky guard flag of the old one goes
	if (a->particleNearWall) h->tileWallValid = false;
	if (a->pos && soft) {
		// the bucket order and the candidate lists survive an upload that moves nobody beyond their guards (a host-driven loop
		// that sends back the positions it received): the displacement trackers decide, exactly as inside a step
		float md = 0.f;
		CUDA_TRY(h, cudaMemcpyAsync(h->h_flags, S.flags, sizeof(int32_t) * NFLAGS, cudaMemcpyDeviceToHost, h->stream));
		CUDA_TRY(h, cudaMemcpyAsync(&md, S.maxDisp2, sizeof(float), cudaMemcpyDeviceToHost, h->stream));
		CUDA_TRY(h, cudaStreamSynchronize(h->stream));
		h->hitValid = false;
		if (h->h_flags[FLAG_UPLOAD_TYPES]) soft = false;
		else h->dispNow = fmax(h->dispNow, sqrt((double)md));
	}
	if (a->pos && !soft) { h->hitValid = false; h->hitForce = true; h->needRebuild = true; }   // new positions: the next fused step re-sorts and rebuilds the lists
	if (a->pos && !soft && h->K.p.domainTypeBC == 1) {   // a new state: nobody is hidden by a periodic wrap of the old one
		CUDA_TRY(h, cudaMemsetAsync(S.wrapHid, 0, (size_t)S.n0, h->stream)); CUDA_TRY(h, cudaMemsetAsync(S.flags + FLAG_WRAP_STEP, 0, sizeof(int32_t), h->stream));
		h->nWrappedCleared = h->nWrapped;
	}
	return PMPS_OK;
}

extern "C" int pmps_upload(pmps_t* h, const PmpsHostArrays* a) { return upload_impl(h, a, false); }
extern "C" int pmps_upload_device(pmps_t* h, const PmpsHostArrays* a) { return upload_impl(h, a, true); }
extern "C" int pmps_upload_values_device(pmps_t* h, const PmpsHostArrays* a) { return upload_impl(h, a, true, true); }

static int download_impl(pmps_handle* h, PmpsHostArrays* a, bool deviceDst) {
	if (!h || !a) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_download: nothing uploaded");
	if (a->numParticlesZero != h->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_download: numParticlesZero mismatch");
	const int n0 = h->numParticlesZero, T = 256, G = cdiv(n0, T);
	DevView& S = h->V; const int32_t* soh = S.slotOfHost;
	// two staging halves so the gather of field k+1 can be queued behind the copy-out of field k
	char* half[2] = { (char*)h->stage, (char*)h->stage + 32 * (size_t)n0 }; int hb = 0;
	auto out = [&](void* dstHost, const void* stg, size_t bytes) { if (deviceDst) return; cudaMemcpyAsync(dstHost, stg, bytes, cudaMemcpyDeviceToHost, h->stream); };
	// device destinations (pmps_download_device): the gather kernels write straight into the caller's arrays
#define STG(dst) (deviceDst ? (char*)(dst) : half[(hb ^= 1) ^ 1])
	if (a->pos || a->particleType || a->PTYPE) {
		char* b = half[hb]; hb ^= 1;
		double* x = a->pos ? (double*)b : nullptr; int32_t* t = a->particleType ? (int32_t*)(b + 24 * (size_t)n0) : nullptr;
		int32_t* pt = a->PTYPE ? (int32_t*)(b + 28 * (size_t)n0) : nullptr;
		if (deviceDst) { x = a->pos; t = a->particleType; pt = a->PTYPE; }
		LAUNCH(h, "download", k_dn_pack, G, T, x, t, pt, S.pos, soh, n0, (const int32_t*)nullptr);
		if (x) out(a->pos, x, 24 * (size_t)n0);
		if (t) out(a->particleType, t, 4 * (size_t)n0);
		if (pt) out(a->PTYPE, pt, 4 * (size_t)n0);
	}
	if (a->vel) { char* b = STG(a->vel); LAUNCH(h, "download", k_dn_pack, G, T, (double*)b, (int32_t*)nullptr, (int32_t*)nullptr, S.vel, soh, n0, (const int32_t*)nullptr); out(a->vel, b, 24 * (size_t)n0); }
	if (a->particleNearWall) { char* b = STG(a->particleNearWall); LAUNCH(h, "download", k_dn_u8, G, T, (uint8_t*)b, S.particleNearWall, soh, n0, (const int32_t*)nullptr); out(a->particleNearWall, b, (size_t)n0); }
#define X(n) if (a->n) { char* b = STG(a->n); LAUNCH(h, "download", k_dn_i32, G, T, (int32_t*)b, S.n, soh, n0, (const int32_t*)nullptr); out(a->n, b, 4 * (size_t)n0); }
	PMPS_I32_FIELDS(X)
#undef X
#define X(n) if (a->n) { char* b = STG(a->n); LAUNCH(h, "download", k_dn_f64, G, T, (double*)b, S.n, soh, n0, (const int32_t*)nullptr); out(a->n, b, 8 * (size_t)n0); }
	PMPS_F64_FIELDS(X)
	if (h->K.p.fluidType == 1) { PMPS_RHEO_FIELDS(X) }
#undef X
#define X(n) if (a->n) { char* b = STG(a->n); LAUNCH(h, "download", k_dn_v3, G, T, (double*)b, S.n, soh, n0, (const int32_t*)nullptr); out(a->n, b, 24 * (size_t)n0); }
	PMPS_V3_FIELDS(X)
#undef X
	if (h->deferSync) return PMPS_OK;   // pmps_step_host: this part rides on the copy stream while the step goes on
	CUDA_TRY(h, cudaMemcpyAsync(h->h_flags, S.flags, sizeof(int32_t) * NFLAGS, cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	h->numParticles = h->h_flags[FLAG_LIVE];
	a->numParticles = h->numParticles;
	return PMPS_OK;
#undef STG
}
// ---- row sets (slab re-partition: only the rows that change hands or places travel; polymps_b200/slab.py) --------------------------
// Gathers / scatters the particles with host indices d_rows[0..n) from / into COMPACT device arrays of n entries each (the arrays of
// `a` that are not NULL).  The upload turns the handle's particle table into a new one: bucket order and lists are re-made before
// the next pass, numParticles becomes the new live count.
extern "C" int pmps_download_rows_device(pmps_t* h, PmpsHostArrays* a, int n, const int32_t* d_rows) {
	if (!h || !a || n < 0 || (n > 0 && !d_rows)) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_download_rows_device: nothing uploaded");
	if (n == 0) return PMPS_OK;
	const int T = 256, G = cdiv(n, T); DevView& S = h->V; const int32_t* soh = S.slotOfHost;
	if (a->pos || a->particleType || a->PTYPE) LAUNCH(h, "download", k_dn_pack, G, T, a->pos, a->particleType, a->PTYPE, S.pos, soh, n, d_rows);
	if (a->vel) LAUNCH(h, "download", k_dn_pack, G, T, a->vel, (int32_t*)nullptr, (int32_t*)nullptr, S.vel, soh, n, d_rows);
	if (a->particleNearWall) LAUNCH(h, "download", k_dn_u8, G, T, a->particleNearWall, S.particleNearWall, soh, n, d_rows);
#define X(f) if (a->f) LAUNCH(h, "download", k_dn_i32, G, T, a->f, S.f, soh, n, d_rows);
	PMPS_I32_FIELDS(X)
#undef X
#define X(f) if (a->f) LAUNCH(h, "download", k_dn_f64, G, T, a->f, S.f, soh, n, d_rows);
	PMPS_F64_FIELDS(X)
	if (h->K.p.fluidType == 1) { PMPS_RHEO_FIELDS(X) }
#undef X
#define X(f) if (a->f) LAUNCH(h, "download", k_dn_v3, G, T, a->f, S.f, soh, n, d_rows);
	PMPS_V3_FIELDS(X)
#undef X
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
extern "C" int pmps_upload_rows_device(pmps_t* h, const PmpsHostArrays* a, int n, const int32_t* d_rows, int numParticles) {
	if (!h || !a || n < 0 || (n > 0 && !d_rows)) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_upload_rows_device: upload a full particle table first");
	if (numParticles < 0 || numParticles > h->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_upload_rows_device: bad numParticles");
	const int T = 256, G = cdiv(n > 0 ? n : 1, T); DevView& S = h->V; const int32_t* soh = S.slotOfHost;
	if (n > 0) {
		if (a->pos || a->particleType || a->PTYPE) LAUNCH(h, "upload", k_up_pack, G, T, (const double*)a->pos, (const int32_t*)a->particleType, (const int32_t*)a->PTYPE, S.pos, soh, n, d_rows);
		if (a->vel) LAUNCH(h, "upload", k_up_pack, G, T, (const double*)a->vel, (const int32_t*)nullptr, (const int32_t*)nullptr, S.vel, soh, n, d_rows);
		if (a->particleNearWall) LAUNCH(h, "upload", k_up_u8, G, T, (const uint8_t*)a->particleNearWall, S.particleNearWall, soh, n, d_rows);
#define X(f) if (a->f) LAUNCH(h, "upload", k_up_i32, G, T, (const int32_t*)a->f, S.f, soh, n, d_rows);
		PMPS_I32_FIELDS(X)
#undef X
#define X(f) if (a->f) LAUNCH(h, "upload", k_up_f64, G, T, (const double*)a->f, S.f, soh, n, d_rows);
		PMPS_F64_FIELDS(X)
		if (h->K.p.fluidType == 1) { PMPS_RHEO_FIELDS(X) }
#undef X
#define X(f) if (a->f) LAUNCH(h, "upload", k_up_v3, G, T, (const double*)a->f, S.f, soh, n, d_rows);
		PMPS_V3_FIELDS(X)
#undef X
	}
	LAUNCH(h, "upload", k_set_live, 1, 1, S, numParticles);
	CUDA_TRY(h, cudaGetLastError());
	h->numParticles = numParticles; h->guardWorst = 0.0; h->tileWallValid = false;
	h->hitValid = false; h->hitForce = true; h->needRebuild = true;
	return PMPS_OK;
}

extern "C" int pmps_download(pmps_t* h, PmpsHostArrays* a) { return download_impl(h, a, false); }
extern "C" int pmps_download_device(pmps_t* h, PmpsHostArrays* a) { return download_impl(h, a, true); }

// One host-driven time step: upload `in`, step, download `out` — with the copies that the step does not wait for riding on a second
// stream.  What a step needs first (positions, velocities, mirror positions, types) goes ahead on the step's stream; press / pndi /
// particleBC are first read by the PND pass (phase 2) and arrive on the copy stream while the viscosity / prediction / collision
// passes run.  On the way back press / pndi / particleBC / mirrorParticlePos are final once the pressure is known (phase 5) and are
// copied out while the gradient, correction and shifting passes run; positions, velocities and the rest follow at the end.
// Same results as pmps_upload + pmps_step(1) + pmps_download (tests/test_gpu_parity.py::test_host_driven_loop_keeps_lists_and_results).
extern "C" int pmps_step_host(pmps_t* h, const PmpsHostArrays* in, PmpsHostArrays* out) {
	if (!h || !in || !out) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (h->slab) return pmps_fail(h, PMPS_E_STATE, "pmps_step_host: slab handles are driven phase by phase");
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_step_host: upload the full particle state once with pmps_upload first");
	if (!h->ioStream) {
		if (cudaStreamCreateWithFlags(&h->ioStream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&h->evIoUp, cudaEventDisableTiming) != cudaSuccess ||
		    cudaEventCreateWithFlags(&h->evIoMid, cudaEventDisableTiming) != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "pmps_step_host: stream / event creation failed");
		void* q = nullptr;
		if (cudaMalloc(&q, 80 * (size_t)h->numParticlesZero) != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "pmps_step_host: second staging buffer (%zu bytes): %s", h->stageBytes, cudaGetErrorString(cudaGetLastError()));
		h->allocs.push_back(q); h->stage2 = q;
	}
	PmpsHostArrays early = *in; early.press = nullptr; early.pndi = nullptr; early.particleBC = nullptr;
	const int ood0 = h->h_flags ? h->h_flags[FLAG_OOD_TOTAL] : 0;
	int rc = upload_impl(h, &early, false); if (rc) return rc;
	const int n0 = h->numParticlesZero, T = 256, G = cdiv(n0, T);
	char* s2 = (char*)h->stage2;
	// late inputs: the copy engine brings them into the second staging buffer on the copy stream; the (tiny) scatter kernels run on the
	// step's stream at the point of first use, so nothing has to squeeze in between the persistent walker kernels
	const bool haveLate = in->press || in->pndi || in->particleBC;
	if (haveLate) {
		if (in->press) { cudaMemcpyAsync(s2, in->press, 8 * (size_t)n0, cudaMemcpyHostToDevice, h->ioStream); }
		if (in->pndi) { cudaMemcpyAsync(s2 + 8 * (size_t)n0, in->pndi, 8 * (size_t)n0, cudaMemcpyHostToDevice, h->ioStream); }
		if (in->particleBC) { cudaMemcpyAsync(s2 + 16 * (size_t)n0, in->particleBC, 4 * (size_t)n0, cudaMemcpyHostToDevice, h->ioStream); }
		cudaEventRecord(h->evIoUp, h->ioStream);
	}
	auto scatter_late = [&]() {
		if (!haveLate) return;
		cudaStreamWaitEvent(h->stream, h->evIoUp, 0);
		DevView& S = h->V; const int32_t* soh = S.slotOfHost;
		if (in->press) LAUNCH(h, "upload", k_up_f64, G, T, (const double*)s2, S.press, soh, n0, (const int32_t*)nullptr);
		if (in->pndi) LAUNCH(h, "upload", k_up_f64, G, T, (const double*)(s2 + 8 * (size_t)n0), S.pndi, soh, n0, (const int32_t*)nullptr);
		if (in->particleBC) LAUNCH(h, "upload", k_up_i32, G, T, (const int32_t*)(s2 + 16 * (size_t)n0), S.particleBC, soh, n0, (const int32_t*)nullptr);
	};
	unsigned m = 0;
	// the first two phases read these fields with the rheology, the prediction gradient or the dynamic collisions; and a step that
	// re-sorts renumbers the slots the scatter writes into: then the scatter goes first
	const PmpsParams& P = h->K.p;
	const bool first = P.fluidType == 1 || P.relaxPress < 1.0 || P.collisionType == 1 || pmps_slab_wants_rebuild(h) == 1;
	if (first) scatter_late();
	for (int ph = 0; ph < 2; ph++) { rc = pmps_step_phase(h, ph, &m); if (rc) return rc; }
	if (!first) scatter_late();
	for (int ph = 2; ph < 6; ph++) { rc = pmps_step_phase(h, ph, &m); if (rc) return rc; }
	// early outputs: gathered on the step's stream, copied out by the copy stream while phases 6-9 run
	PmpsHostArrays mid, rest = *out; memset(&mid, 0, sizeof mid); mid.numParticles = out->numParticles; mid.numParticlesZero = out->numParticlesZero;
	mid.press = out->press; mid.pndi = out->pndi; mid.particleBC = out->particleBC; mid.mirrorParticlePos = out->mirrorParticlePos;
	const bool haveMid = mid.press || mid.pndi || mid.particleBC || mid.mirrorParticlePos;
	if (haveMid) {
		DevView& S = h->V; const int32_t* soh = S.slotOfHost;
		double* gp = (double*)s2; double* gn = (double*)(s2 + 8 * (size_t)n0); int32_t* gb = (int32_t*)(s2 + 16 * (size_t)n0); double* gm = (double*)(s2 + 24 * (size_t)n0);
		if (mid.press) LAUNCH(h, "download", k_dn_f64, G, T, gp, S.press, soh, n0, (const int32_t*)nullptr);
		if (mid.pndi) LAUNCH(h, "download", k_dn_f64, G, T, gn, S.pndi, soh, n0, (const int32_t*)nullptr);
		if (mid.particleBC) LAUNCH(h, "download", k_dn_i32, G, T, gb, S.particleBC, soh, n0, (const int32_t*)nullptr);
		if (mid.mirrorParticlePos) LAUNCH(h, "download", k_dn_v3, G, T, gm, S.mirrorParticlePos, soh, n0, (const int32_t*)nullptr);
		cudaEventRecord(h->evIoMid, h->stream); cudaStreamWaitEvent(h->ioStream, h->evIoMid, 0);
		if (mid.press) { cudaMemcpyAsync(mid.press, gp, 8 * (size_t)n0, cudaMemcpyDeviceToHost, h->ioStream); }
		if (mid.pndi) { cudaMemcpyAsync(mid.pndi, gn, 8 * (size_t)n0, cudaMemcpyDeviceToHost, h->ioStream); }
		if (mid.particleBC) { cudaMemcpyAsync(mid.particleBC, gb, 4 * (size_t)n0, cudaMemcpyDeviceToHost, h->ioStream); }
		if (mid.mirrorParticlePos) { cudaMemcpyAsync(mid.mirrorParticlePos, gm, 24 * (size_t)n0, cudaMemcpyDeviceToHost, h->ioStream); }
		rest.press = nullptr; rest.pndi = nullptr; rest.particleBC = nullptr; rest.mirrorParticlePos = nullptr;
	}
	rc = pmps_step_phase(h, 6, &m); if (rc) return rc;
	// polygon walls: the velocities are final after the correction (phase 6; with particle walls the dummy velocities change in phase 9)
	const bool velEarly = h->K.poly && rest.vel;
	double* velOut = rest.vel;
	if (velEarly) {
		double* gv = (double*)(s2 + 48 * (size_t)n0);
		LAUNCH(h, "download", k_dn_pack, G, T, gv, (int32_t*)nullptr, (int32_t*)nullptr, h->V.vel, h->V.slotOfHost, n0, (const int32_t*)nullptr);
		cudaEventRecord(h->evIoUp, h->stream); cudaStreamWaitEvent(h->ioStream, h->evIoUp, 0); cudaMemcpyAsync(velOut, gv, 24 * (size_t)n0, cudaMemcpyDeviceToHost, h->ioStream);
		rest.vel = nullptr; mid.vel = velOut;
	}
	for (int ph = 7; ph < pmps_step_phases(); ph++) { rc = pmps_step_phase(h, ph, &m); if (rc) return rc; }
	rc = pmps_step_end_host(h); if (rc) return rc;
	rc = download_impl(h, &rest, false); if (rc) return rc;   // synchronises the step's stream, reads the flags
	if (cudaStreamSynchronize(h->ioStream) != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "pmps_step_host: copy stream: %s", cudaGetErrorString(cudaGetLastError()));
	if ((haveMid || velEarly) && h->h_flags[FLAG_OOD_TOTAL] != ood0) {   // a particle left the domain after the early copy: its rows changed, fetch them again
		rc = download_impl(h, &mid, false); if (rc) return rc;
	}
	out->numParticles = h->numParticles;
	return PMPS_OK;
}
