// pmps_grid.cu — ingestion of the reference's particle file straight into the device layout.
// Replaces MpsInputOutput::readMpsParticleFile (src/MpsInputOutput.cpp:645-857) for hosts that do not want the reference's 585 bytes
// per particle of host arrays (200 M particles = 117 GB): the .grid text — "0", N, then N rows "type x y z vx vy vz press pressAverage"
// (README.md:65-74) — is read in chunks, parsed by a few host threads (strtol / strtod: correctly rounded, the same values fscanf's
// %d / %lf give), and each chunk goes to the device as four small arrays; a kernel builds the 32-byte position / velocity records
// and the initial values of the other live fields exactly as the reference sets them: everything zero, distParticleWall2 = 10e8 lo,
// RHO / PTYPE / MEU from the fluid type (a sediment id 1 becomes fluid of phase 2 in a non-Newtonian run, :800-857).
// Host order = file order, so slot = host index until the first bucket rebuild.
#include "pmps_device.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

__global__ void k_grid_rows(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int first, int n, const int32_t* __restrict__ type,
                            const double* __restrict__ xyz, const double* __restrict__ vel, const double* __restrict__ press) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n) return;
	const int i = first + t;   // host index = slot (identity permutation of a fresh handle)
	int ty = type[t], pty = 1; double rho = K.p.DNS_FL1, meu = K.p.KNM_VS1 * K.p.DNS_FL1;
	if (K.p.fluidType == 1 && ty == 1) { ty = K.p.fluid; pty = 2; rho = K.p.DNS_FL2; meu = K.p.KNM_VS2 * K.p.DNS_FL2; }
	double4 p; p.x = xyz[3 * (size_t)t]; p.y = xyz[3 * (size_t)t + 1]; p.z = xyz[3 * (size_t)t + 2]; p.w = pk_make(ty, pty);
	double4 v; v.x = vel[3 * (size_t)t]; v.y = vel[3 * (size_t)t + 1]; v.z = vel[3 * (size_t)t + 2]; v.w = 0.0;
	S.pos[i] = p; S.vel[i] = v; S.press[i] = press[t];
	S.RHO[i] = rho; S.MEU[i] = meu; S.distParticleWall2[i] = K.big;
}

namespace {
struct Chunk { std::vector<int32_t> type; std::vector<double> xyz, vel, press; };
// parses rows [r0, r1) of the text between p and end (p at the start of a row); returns the position after the last row, or NULL
const char* parse_rows(const char* p, const char* end, size_t r0, size_t r1, Chunk& C) {
	for (size_t r = r0; r < r1; r++) {
		char* q = nullptr;
		const long ty = strtol(p, &q, 10); if (q == p) return nullptr; p = q;
		double b[8];
		for (int k = 0; k < 8; k++) { b[k] = strtod(p, &q); if (q == p) return nullptr; p = q; }
		if (p > end) return nullptr;
		C.type[r] = (int32_t)ty;
		C.xyz[3 * r] = b[0]; C.xyz[3 * r + 1] = b[1]; C.xyz[3 * r + 2] = b[2];
		C.vel[3 * r] = b[3]; C.vel[3 * r + 1] = b[4]; C.vel[3 * r + 2] = b[5];
		C.press[r] = b[6];
	}
	return p;
}
}  // namespace

extern "C" int pmps_load_grid(pmps_t* h, const char* path, int* numParticlesOut) {
	if (!h || !path) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_load_grid: the handle already holds a particle state");
	FILE* fp = fopen(path, "rb");
	if (!fp) return pmps_fail(h, PMPS_E_ARG, "pmps_load_grid: cannot open %s", path);
	long zero = 0, n = 0;
	if (fscanf(fp, "%ld %ld", &zero, &n) != 2 || n < 1 || n > 2000000000l) { fclose(fp); return pmps_fail(h, PMPS_E_ARG, "pmps_load_grid: %s: bad header", path); }
	{ const int rc = pmps_ensure_capacity(h, (int)n); if (rc) { fclose(fp); return rc; } }
	DevView& S = h->V;
	const size_t rowsPerChunk = 1u << 21;   // 2 M rows: ~250 MB of text, 136 MB of parsed arrays
	const unsigned nThreads = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
	std::vector<char> text; std::string carry;
	Chunk C; C.type.resize(rowsPerChunk); C.xyz.resize(3 * rowsPerChunk); C.vel.resize(3 * rowsPerChunk); C.press.resize(rowsPerChunk);
	int32_t* dT = nullptr; double *dX = nullptr, *dV = nullptr, *dP = nullptr;
	if (cudaMalloc(&dT, 4 * rowsPerChunk) != cudaSuccess || cudaMalloc(&dX, 24 * rowsPerChunk) != cudaSuccess || cudaMalloc(&dV, 24 * rowsPerChunk) != cudaSuccess || cudaMalloc(&dP, 8 * rowsPerChunk) != cudaSuccess) {
		fclose(fp); return pmps_fail(h, PMPS_E_CUDA, "pmps_load_grid: staging allocation failed");
	}
	size_t done = 0; int rc = PMPS_OK;
	const size_t textBytes = 256u << 20;
	while (done < (size_t)n && rc == PMPS_OK) {
		// one chunk of text, cut at the last line end (a row never spans two chunks: the tail is carried over)
		text.assign(carry.begin(), carry.end()); carry.clear();
		const size_t have = text.size(); text.resize(have + textBytes + 1);
		const size_t got = fread(text.data() + have, 1, textBytes, fp);
		size_t len = have + got; text[len] = 0;
		if (got == textBytes) { size_t cut = len; while (cut > 0 && text[cut - 1] != '\n') cut--; carry.assign(text.data() + cut, text.data() + len); len = cut; text[len] = 0; }
		// row starts: one pass over the newlines (a row = one line in every file the reference's tools write; blank lines are skipped)
		std::vector<size_t> start; start.reserve(rowsPerChunk + 1);
		{ size_t p = 0; while (p < len) { while (p < len && (text[p] == '\n' || text[p] == '\r' || text[p] == ' ' || text[p] == '\t')) p++; if (p >= len) break; start.push_back(p); while (p < len && text[p] != '\n') p++; } }
		size_t rows = std::min(start.size(), (size_t)n - done);
		for (size_t r0 = 0; r0 < rows && rc == PMPS_OK; r0 += rowsPerChunk) {
			const size_t r1 = std::min(rows, r0 + rowsPerChunk), m = r1 - r0;
			std::vector<std::thread> th; std::vector<int> ok(nThreads, 1);
			for (unsigned t = 0; t < nThreads; t++) th.emplace_back([&, t]() {
				const size_t a = r0 + m * t / nThreads, b = r0 + m * (t + 1) / nThreads;
				for (size_t r = a; r < b; r++) if (!parse_rows(text.data() + start[r], text.data() + len, r - r0, r - r0 + 1, C)) { ok[t] = 0; return; }
			});
			for (auto& x : th) x.join();
			for (unsigned t = 0; t < nThreads; t++) if (!ok[t]) rc = pmps_fail(h, PMPS_E_ARG, "pmps_load_grid: %s: malformed row near particle %zu", path, done + r0);
			if (rc) break;
			cudaMemcpyAsync(dT, C.type.data(), 4 * m, cudaMemcpyHostToDevice, h->stream); cudaMemcpyAsync(dX, C.xyz.data(), 24 * m, cudaMemcpyHostToDevice, h->stream);
			cudaMemcpyAsync(dV, C.vel.data(), 24 * m, cudaMemcpyHostToDevice, h->stream); cudaMemcpyAsync(dP, C.press.data(), 8 * m, cudaMemcpyHostToDevice, h->stream);
			k_grid_rows<<<cdiv((long long)m, 256), 256, 0, h->stream>>>(S, h->K, (int)(done + r0), (int)m, dT, dX, dV, dP); h->launches++;
			if (cudaStreamSynchronize(h->stream) != cudaSuccess) rc = pmps_fail(h, PMPS_E_CUDA, "pmps_load_grid: %s", cudaGetErrorString(cudaGetLastError()));
		}
		done += rows;
		if (got < textBytes && done < (size_t)n) { rc = pmps_fail(h, PMPS_E_ARG, "pmps_load_grid: %s ends after %zu of %ld particles", path, done, n); break; }
	}
	fclose(fp); cudaFree(dT); cudaFree(dX); cudaFree(dV); cudaFree(dP);
	if (rc) return rc;
	int32_t live = (int32_t)n;
	cudaMemcpyAsync(&S.flags[FLAG_LIVE], &live, sizeof live, cudaMemcpyHostToDevice, h->stream); cudaStreamSynchronize(h->stream);
	h->numParticles = (int)n; h->haveState = true; h->guardWorst = 0.0;
	h->hitValid = false; h->hitForce = true; h->needRebuild = true;
	if (numParticlesOut) *numParticlesOut = (int)n;
	return PMPS_OK;
}
