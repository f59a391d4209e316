// pmps_observe.cu — integrated observables on the device (north_star: "dam-break surge-front position and
// pressure-probe history").  The reference has no device and samples on the host: its pressure sensors average the pressure of the
// particles inside a window (src/MpsInputOutput.cpp:1937-2229), its output steps write every array (:861-1875).  Here one pass over
// the position records gives the extent of a particle type (surge front = largest x of the fluid) and the windowed pressure
// statistics, so a run can record its history every step without a single particle array leaving the GPU.
// Two-stage reductions in a fixed order: the sums are reproducible run to run.
#include "pmps_device.cuh"
#include <cfloat>
#include <mutex>

#define OBS_T 256
struct ObsPartial { double mn[3], mx[3], psum, pmax; long long n, nbox; };

__device__ __forceinline__ void obs_merge(ObsPartial& a, const ObsPartial& b) {
	for (int k = 0; k < 3; k++) { a.mn[k] = fmin(a.mn[k], b.mn[k]); a.mx[k] = fmax(a.mx[k], b.mx[k]); }
	a.psum += b.psum; a.pmax = fmax(a.pmax, b.pmax); a.n += b.n; a.nbox += b.nbox;
}
__device__ __forceinline__ ObsPartial obs_shfl(const ObsPartial& a, int o) {
	ObsPartial b;
	for (int k = 0; k < 3; k++) { b.mn[k] = __shfl_down_sync(0xffffffffu, a.mn[k], o); b.mx[k] = __shfl_down_sync(0xffffffffu, a.mx[k], o); }
	b.psum = __shfl_down_sync(0xffffffffu, a.psum, o); b.pmax = __shfl_down_sync(0xffffffffu, a.pmax, o);
	b.n = __shfl_down_sync(0xffffffffu, a.n, o); b.nbox = __shfl_down_sync(0xffffffffu, a.nbox, o);
	return b;
}
__device__ __forceinline__ ObsPartial obs_empty() {
	ObsPartial p; for (int k = 0; k < 3; k++) { p.mn[k] = DBL_MAX; p.mx[k] = -DBL_MAX; }
	p.psum = 0.0; p.pmax = -DBL_MAX; p.n = 0; p.nbox = 0; return p;
}
__device__ __forceinline__ ObsPartial obs_block_reduce(ObsPartial p, ObsPartial* sh) {   // fixed tree: lanes, then warps in order
	for (int o = 16; o > 0; o >>= 1) { const ObsPartial q = obs_shfl(p, o); obs_merge(p, q); }
	if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = p;
	__syncthreads();
	if (threadIdx.x == 0) { for (int w = 1; w < OBS_T / 32; w++) obs_merge(p, sh[w]); }
	return p;   // valid in thread 0
}
// stage 1: every block walks a contiguous stretch of slots (owned particles only on a slab rank)
__global__ void __launch_bounds__(OBS_T) k_observe(const __grid_constant__ DevView S, int type, double lx, double ly, double lz, double hx, double hy, double hz, int useBox, ObsPartial* out) {
	__shared__ ObsPartial sh[OBS_T / 32];
	ObsPartial p = obs_empty();
	const int per = (S.n + gridDim.x - 1) / gridDim.x, i0 = blockIdx.x * per, i1 = min(S.n, i0 + per);
	for (int i = i0 + threadIdx.x; i < i1; i += OBS_T) {
		const double4 q = ld256(S.pos + i);
		if (pk_type(q) != type || !owned_slot(S, i)) continue;
		p.n++;
		p.mn[0] = fmin(p.mn[0], q.x); p.mn[1] = fmin(p.mn[1], q.y); p.mn[2] = fmin(p.mn[2], q.z);
		p.mx[0] = fmax(p.mx[0], q.x); p.mx[1] = fmax(p.mx[1], q.y); p.mx[2] = fmax(p.mx[2], q.z);
		if (useBox && q.x >= lx && q.x <= hx && q.y >= ly && q.y <= hy && q.z >= lz && q.z <= hz) { const double pr = S.press[i]; p.nbox++; p.psum += pr; p.pmax = fmax(p.pmax, pr); }
	}
	p = obs_block_reduce(p, sh);
	if (threadIdx.x == 0) out[blockIdx.x] = p;
}
__global__ void __launch_bounds__(OBS_T) k_observe_final(ObsPartial* parts, int n) {
	__shared__ ObsPartial sh[OBS_T / 32];
	ObsPartial p = obs_empty();
	for (int k = threadIdx.x; k < n; k += OBS_T) obs_merge(p, parts[k]);
	p = obs_block_reduce(p, sh);
	if (threadIdx.x == 0) parts[0] = p;
}

static int observe(pmps_handle* h, int type, const double* lo, const double* hi, ObsPartial* res) {
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_observe: no particle state uploaded");
	const int grid = h->smCount * 4;
	if (h->stageBytes < sizeof(ObsPartial) * (size_t)grid) return pmps_fail(h, PMPS_E_STATE, "pmps_observe: staging area too small");
	ObsPartial* parts = (ObsPartial*)h->stage;
	LAUNCH(h, "observe", k_observe, grid, OBS_T, h->V, type, lo ? lo[0] : 0.0, lo ? lo[1] : 0.0, lo ? lo[2] : 0.0, hi ? hi[0] : 0.0, hi ? hi[1] : 0.0, hi ? hi[2] : 0.0, lo && hi ? 1 : 0, parts);
	LAUNCH(h, "observe", k_observe_final, 1, OBS_T, parts, grid);
	CUDA_TRY(h, cudaMemcpyAsync(res, parts, sizeof(ObsPartial), cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	return PMPS_OK;
}
extern "C" int pmps_observe_extent(pmps_t* h, int particleType, double mn[3], double mx[3], long long* count) {
	if (!h || !mn || !mx) return PMPS_E_ARG;
	ObsPartial r; const int rc = observe(h, particleType, nullptr, nullptr, &r); if (rc) return rc;
	for (int k = 0; k < 3; k++) { mn[k] = r.mn[k]; mx[k] = r.mx[k]; }
	if (count) *count = r.n;
	return PMPS_OK;
}
extern "C" int pmps_observe_box(pmps_t* h, int particleType, const double lo[3], const double hi[3], long long* count, double* pressSum, double* pressMax) {
	if (!h || !lo || !hi) return PMPS_E_ARG;
	ObsPartial r; const int rc = observe(h, particleType, lo, hi, &r); if (rc) return rc;
	if (count) *count = r.nbox;
	if (pressSum) *pressSum = r.psum;
	if (pressMax) *pressMax = r.nbox ? r.pmax : 0.0;
	return PMPS_OK;
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// pmps_observe_extent_static: the extent reduction with its per-block partial records in a file-scope device array instead of
// the handle's cudaMalloc'ed staging area.  No counterpart in the reference; added for
// callers that may not allocate (graph capture, tight serving loops).  The call makes no cudaMalloc / cudaFree and uses no
// allocating library routine.  Limits and rules:
//   * the array holds PMPS_OBS_STATIC_MAX partial records (1024 x 80 bytes = 80 KiB); every GPU that loads the library keeps it
//     until exit.  One record per block, four blocks per SM: a device that would need more records is refused with PMPS_E_ARG;
//   * every call on a device shares the one array.  Stage 1 overwrites each record it later reads, so nothing has to start at
//     zero and the previous call's contents are harmless, but two calls that overlap on the same GPU (different handles or
//     streams) race on it: serialise them;
//   * in host code the array's name is only a stand-in; its device address is looked up once per device, with that device
//     current, through cudaGetSymbolAddress and handed to the kernels.
#define PMPS_OBS_STATIC_MAX 1024
#define PMPS_OBS_MAX_DEVICES 64
static __device__ __align__(256) ObsPartial g_obsScratch[PMPS_OBS_STATIC_MAX];
static ObsPartial* g_obsScratchAddr[PMPS_OBS_MAX_DEVICES];   // per device ordinal; written once under the lock below
static std::mutex g_obsScratchLock;

static int obs_static_scratch(pmps_handle* h, ObsPartial** out) {   // h->device is current
	if (h->device < 0 || h->device >= PMPS_OBS_MAX_DEVICES) return pmps_fail(h, PMPS_E_ARG, "pmps_observe_extent_static: device ordinal %d beyond %d", h->device, PMPS_OBS_MAX_DEVICES - 1);
	std::lock_guard<std::mutex> g(g_obsScratchLock);
	if (!g_obsScratchAddr[h->device]) {
		ObsPartial* p = nullptr;
		CUDA_TRY(h, cudaGetSymbolAddress((void**)&p, g_obsScratch));
		if (!p) return pmps_fail(h, PMPS_E_CUDA, "pmps_observe_extent_static: cudaGetSymbolAddress returned no address");
		g_obsScratchAddr[h->device] = p;
	}
	*out = g_obsScratchAddr[h->device];
	return PMPS_OK;
}

extern "C" int pmps_observe_extent_static(pmps_t* h, int particleType, double mn[3], double mx[3], long long* count) {
	if (!h || !mn || !mx) return PMPS_E_ARG;
	CUDA_TRY(h, cudaSetDevice(h->device));
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_observe_extent_static: no particle state uploaded");
	const int grid = h->smCount * 4;   // same partition as pmps_observe_extent
	if (grid > PMPS_OBS_STATIC_MAX)
		return pmps_fail(h, PMPS_E_ARG, "pmps_observe_extent_static: %d partial records needed (4 per SM, %d SMs), the static scratch holds %d", grid, h->smCount, PMPS_OBS_STATIC_MAX);
	ObsPartial* parts = nullptr;
	const int rc = obs_static_scratch(h, &parts); if (rc) return rc;
	k_observe<<<grid, OBS_T, 0, h->stream>>>(h->V, particleType, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0, parts);
	k_observe_final<<<1, OBS_T, 0, h->stream>>>(parts, grid);
	h->launches += 2;
	CUDA_TRY(h, cudaGetLastError());
	ObsPartial r;
	CUDA_TRY(h, cudaMemcpyAsync(&r, parts, sizeof(ObsPartial), cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	for (int k = 0; k < 3; k++) { mn[k] = r.mn[k]; mx[k] = r.mx[k]; }
	if (count) *count = r.n;
	return PMPS_OK;
}
