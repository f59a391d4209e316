// pmps_slab.cu — the device half of the slab decomposition over several GPUs.
//
// One process per GPU owns the particles of one x-slab of the domain and keeps HALO COPIES of the neighbouring ranks'
// particles within one search radius (bucketSide) of its cuts.  Halo copies sit at host indices [nOwned, numParticles):
// they are sorted, staged and visited like every other particle, so the neighbour sums of the owned particles are
// complete, and whatever the passes compute FOR them is overwritten by the owner's values before anybody reads it:
// after every phase of the step (pmps_step_phase) the caller packs the fields that phase produced for the particles its
// neighbours hold copies of (pmps_halo_pack), exchanges the buffers (NCCL send/recv, polymps_b200/slab.py) and unpacks
// what it received onto its own copies (pmps_halo_unpack).  Buffers are plane-major: plane p of particle t at [p*n + t].
// The Poisson solve of the incompressible scheme (pmps_ppe.cu) reaches the host through pmps_slab_set_hooks for its all-reduces
// and the refresh of the search direction.
#include "pmps_device.cuh"

static __host__ __device__ inline int halo_planes(unsigned f) {
	return ((f & PMPS_HF_POS) ? 4 : 0) + ((f & PMPS_HF_VEL) ? 3 : 0) + ((f & PMPS_HF_PRESS) ? 1 : 0) + ((f & PMPS_HF_PNDI) ? 1 : 0)
	     + ((f & PMPS_HF_CONC) ? 1 : 0) + ((f & PMPS_HF_PNDKI) ? 1 : 0) + ((f & PMPS_HF_BC) ? 1 : 0) + ((f & PMPS_HF_PPE_X) ? 1 : 0) + ((f & PMPS_HF_PPE_P) ? 1 : 0);
}
extern "C" int pmps_halo_planes(unsigned fields) { return halo_planes(fields); }

__global__ void k_halo_pack(const __grid_constant__ DevView S, unsigned f, const int32_t* __restrict__ hostIndex, int n, double* __restrict__ out) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n) return;
	const int s = S.slotOfHost[hostIndex[t]];
	double* o = out + t; const size_t st = (size_t)n;
	if (f & PMPS_HF_POS) { const double4 p = S.pos[s]; o[0] = p.x; o[st] = p.y; o[2 * st] = p.z; o[3 * st] = p.w; o += 4 * st; }
	if (f & PMPS_HF_VEL) { const double4 v = S.vel[s]; o[0] = v.x; o[st] = v.y; o[2 * st] = v.z; o += 3 * st; }
	if (f & PMPS_HF_PRESS) { o[0] = S.press[s]; o += st; }
	if (f & PMPS_HF_PNDI) { o[0] = S.pndi[s]; o += st; }
	if (f & PMPS_HF_CONC) { o[0] = S.concentration[s]; o += st; }
	if (f & PMPS_HF_PNDKI) { o[0] = S.pndki[s]; o += st; }
	if (f & PMPS_HF_BC) { o[0] = (double)S.particleBC[s]; o += st; }
	if (f & PMPS_HF_PPE_X) { o[0] = S.pressurePPE[s]; o += st; }
	if (f & PMPS_HF_PPE_P) { o[0] = S.ppeP[s]; o += st; }
}
__global__ void k_halo_unpack(const __grid_constant__ DevView S, const __grid_constant__ KParams K, unsigned f, int first, int n, const double* __restrict__ in) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n) return;
	const int s = S.slotOfHost[first + t];
	const double* o = in + t; const size_t st = (size_t)n;
	if (f & PMPS_HF_POS) {
		double4 p; p.x = o[0]; p.y = o[st]; p.z = o[2 * st]; p.w = o[3 * st];
		S.pos[s] = p; o += 4 * st;
		track_hit_disp(S, K, s, p);   // the copy's TRUE position: does the hit list still cover it?

	}
	if (f & PMPS_HF_VEL) { double4 v = S.vel[s]; v.x = o[0]; v.y = o[st]; v.z = o[2 * st]; S.vel[s] = v; o += 3 * st; }
	if (f & PMPS_HF_PRESS) { S.press[s] = o[0]; o += st; }
	if (f & PMPS_HF_PNDI) { S.pndi[s] = o[0]; o += st; }
	if (f & PMPS_HF_CONC) { S.concentration[s] = o[0]; o += st; }
	if (f & PMPS_HF_PNDKI) { S.pndki[s] = o[0]; o += st; }
	if (f & PMPS_HF_BC) { S.particleBC[s] = (int)o[0]; o += st; }
	if (f & PMPS_HF_PPE_X) { S.pressurePPE[s] = o[0]; o += st; }
	if (f & PMPS_HF_PPE_P) { S.ppeP[s] = o[0]; o += st; }
}

extern "C" int pmps_slab_begin(pmps_t* h, int nOwned) {
	if (!h) return PMPS_E_ARG;
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_slab_begin: upload the particle table first");
	if (h->K.p.fluidType == 1) return pmps_fail(h, PMPS_E_UNSUPPORTED, "pmps_slab_begin: the non-Newtonian rheology (column maps, Cv / MEU / RHO halo fields) is single-GPU so far");
	if (h->K.p.domainTypeBC == 1) return pmps_fail(h, PMPS_E_UNSUPPORTED, "pmps_slab_begin: periodic domains are single-GPU so far (the slab driver has no wrap-around halo)");
	if (nOwned < 0 || nOwned > h->numParticles) return pmps_fail(h, PMPS_E_ARG, "pmps_slab_begin: %d owned of %d particles", nOwned, h->numParticles);
	const PmpsParams& P = h->K.p;
	if ((P.mpsType == 2 || P.mpsType == 3) && !h->slabSum) return pmps_fail(h, PMPS_E_STATE, "pmps_slab_begin: the incompressible scheme needs pmps_slab_set_hooks first (all-reduce / halo refresh inside the Poisson solve)");
	h->V.nOwn = nOwned; h->slab = true;
	return PMPS_OK;
}
extern "C" int pmps_slab_set_hooks(pmps_t* h, double* d_workspace8, pmps_sum_fn sum, pmps_refresh_fn refresh, void* user) {
	if (!h || !d_workspace8 || !sum || !refresh) return PMPS_E_ARG;
	h->slabWork = d_workspace8; h->slabSum = sum; h->slabRefresh = refresh; h->slabUser = user;
	return PMPS_OK;
}
extern "C" int pmps_halo_pack(pmps_t* h, unsigned fields, const int32_t* d_hostIndex, int n, double* d_out) {
	if (!h || n < 0 || (n > 0 && (!d_hostIndex || !d_out))) return PMPS_E_ARG;
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_halo_pack: no bucket order yet");
	if (n == 0 || fields == 0) return PMPS_OK;
	if ((fields & PMPS_HF_PPE_P) && !h->V.ppeP) return pmps_fail(h, PMPS_E_STATE, "pmps_halo_pack: no Poisson solve in progress");
	cudaSetDevice(h->device);
	LAUNCH(h, "halo.pack", k_halo_pack, cdiv(n, 256), 256, h->V, fields, d_hostIndex, n, d_out);
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
extern "C" int pmps_halo_unpack(pmps_t* h, unsigned fields, int firstHostIndex, int n, const double* d_in) {
	if (!h || n < 0 || firstHostIndex < 0 || (n > 0 && !d_in)) return PMPS_E_ARG;
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_halo_unpack: no bucket order yet");
	if ((long long)firstHostIndex + n > h->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_halo_unpack: [%d, %d) exceeds the capacity %d", firstHostIndex, firstHostIndex + n, h->numParticlesZero);
	if (firstHostIndex < h->V.nOwn && n > 0) return pmps_fail(h, PMPS_E_ARG, "pmps_halo_unpack: host index %d is an owned particle (nOwned %d)", firstHostIndex, h->V.nOwn);
	if (n == 0 || fields == 0) return PMPS_OK;
	cudaSetDevice(h->device);
	LAUNCH(h, "halo.unpack", k_halo_unpack, cdiv(n, 256), 256, h->V, h->K, fields, firstHostIndex, n, d_in);
	CUDA_TRY(h, cudaGetLastError());
	if (fields & PMPS_HF_POS) h->hitValid = false;   // checked (and rebuilt if needed) before the next full pass
	return PMPS_OK;
}

// step scalars: {max squared displacement since the list build, max squared shift of this step, velMax^2, the displacement the step
// before ended with}, FOUR doubles in
// DEVICE memory so that the caller can all-reduce (MAX) them in place over the ranks before pmps_slab_apply reads them
__global__ void k_slab_collect(const __grid_constant__ DevView S, double* out) {
	out[0] = (double)S.maxDisp2[0]; out[1] = (double)S.maxDisp2[1]; out[2] = __longlong_as_double((long long)S.velMax2Bits[1]);
	out[3] = (double)S.maxDisp2[2];   // the displacement the step before ended with (pmps_slab_mid_check)
}
__global__ void k_slab_set_velmax(const __grid_constant__ DevView S, const double* in) {
	const unsigned long long b = (unsigned long long)__double_as_longlong(in[2]); S.velMax2Bits[0] = b; S.velMax2Bits[1] = b;
	S.maxDisp2[2] = S.maxDisp2[0];   // the displacement this step ends with: the next mid-step check derives the size of a step's small moves from it
}
extern "C" int pmps_slab_collect(pmps_t* h, double* d_out4) {
	double* const d_out3 = d_out4;
	if (!h || !d_out3) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	LAUNCH(h, "slab.collect", k_slab_collect, 1, 1, h->V, d_out3);
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
extern "C" int pmps_slab_apply(pmps_t* h, const double* d_in3) {
	if (!h || !d_in3) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	LAUNCH(h, "slab.set_velmax", k_slab_set_velmax, 1, 1, h->V, d_in3);
	double v[3] = { 0.0, 0.0, 0.0 };
	CUDA_TRY(h, cudaMemcpyAsync(v, d_in3, sizeof v, cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	h->velMax = sqrt(v[2]);
	return pmps_step_end_apply(h, v[0], v[1]);
}
