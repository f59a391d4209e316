// pmps_vtu.cu — device-fed VTU output: the .vtu file of MpsInputOutput::writeVtuAscii /
// writeVtuAsciiFreeSurface (reference src/MpsInputOutput.cpp:1313-1875), byte for byte, straight from the device state.
//
// The reference formats every number with its own fprintf call from one thread: 25 - 45 calls per particle, tens of seconds per
// output step at 10 M particles — as long as a few hundred time steps of the device path.  Here only the arrays the file holds
// come back from the device (pmps_download, host order), and the text is produced by all host cores: the particles are cut into
// blocks, every thread formats whole blocks into its own buffer with an exact fixed-point formatter (same digits as printf's
// "%f": round-half-even of the EXACT binary value, checked against printf in tests/test_vtu.py), and the blocks are written in order.
// Host code only: no kernel of its own.
#include "pmps_internal.h"
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

// ---- "%f" (6 decimals) of a double, as glibc prints it -----------------------------------------------------------------------
// |x| = ip + frac with ip an integer < 2^52 and frac in [0, 1), both exact.  The six decimals are round-half-even(frac * 1e6) of the
// EXACT product: hi = fl(frac * 1e6), lo = fma(frac, 1e6, -hi) is its exact error (1e6 is exact), r0 = floor(hi) and hi - r0 are
// exact, so  frac * 1e6 - r0 - 1/2 = (hi - r0 - 1/2) + lo  is the sum of two exact doubles whose sign is decided by one exact
// comparison.  Values too large for that go through snprintf.
static inline char* put_uint(char* p, unsigned long long v) {
	char tmp[24]; int n = 0;
	do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
	while (n) *p++ = tmp[--n];
	return p;
}
static inline char* put_int(char* p, int v) {
	unsigned long long u = v < 0 ? (unsigned long long)(-(long long)v) : (unsigned long long)v;
	if (v < 0) *p++ = '-';
	return put_uint(p, u);
}
static char* put_f6(char* p, double x) {
	if (!(fabs(x) < 4.0e15)) { p += snprintf(p, 400, "%f", x); return p; }   // huge, inf, nan: the library's own conversion
	if (std::signbit(x)) { *p++ = '-'; x = -x; }
	double ipd = floor(x); const double frac = x - ipd;   // exact
	const double hi = frac * 1.0e6, lo = fma(frac, 1.0e6, -hi);
	double r0 = floor(hi), dh = hi - r0;                   // exact, dh in [0, 1)
	if (dh == 0.0 && lo < 0.0) { r0 -= 1.0; dh = 1.0; }    // the exact product lies just below the integer hi
	unsigned long long r = (unsigned long long)r0;
	const double t = 0.5 - dh;                             // exact; frac * 1e6 - r0 > 1/2  <=>  lo > t
	if (lo > t || (lo == t && (r & 1ull))) r++;
	unsigned long long ip = (unsigned long long)ipd;
	if (r >= 1000000ull) { r -= 1000000ull; ip++; }
	p = put_uint(p, ip);
	*p++ = '.';
	for (int k = 5; k >= 0; k--) { p[k] = (char)('0' + r % 10); r /= 10; }
	return p + 6;
}
extern "C" int pmps_debug_format_f6(const double* v, int n, char* out, size_t cap) {   // test hook: "%f " of every value
	if (!v || !out || n < 0) return PMPS_E_ARG;
	char* p = out;
	for (int i = 0; i < n; i++) { if ((size_t)(p - out) + 420 > cap) return PMPS_E_OVERFLOW; p = put_f6(p, v[i]); *p++ = ' '; }
	*p = 0;
	return (int)(p - out);
}

// ---- the arrays of the file ----------------------------------------------------------------------------------------------------
namespace {
enum Kind { K_POS3, K_F3, K_F1, K_I1, K_SEQ0, K_SEQ1, K_ONE };   // positions ("%lf" of the double), (float) vectors / scalars, ints, 0.., 1.., "1 "
struct Arr { const char* open; Kind kind; const double* d; const int32_t* i; };

struct Writer {
	FILE* fp; int n; const std::vector<int32_t>* sel;   // sel: the particles of the free-surface variant (ascending), or null = all
	int nThreads;
	int count() const { return sel ? (int)sel->size() : n; }
	// formats entries [a, b) of the selection into buf
	void format(const Arr& A, int a, int b, std::string& buf) const {
		buf.resize((size_t)(b - a) * (A.kind == K_POS3 || A.kind == K_F3 ? 3 * 26 : 26) + 16);
		char* p = &buf[0];
		for (int e = a; e < b; e++) {
			const int i = sel ? (*sel)[e] : e;
			switch (A.kind) {
				case K_POS3: for (int k = 0; k < 3; k++) { p = put_f6(p, A.d[3 * (size_t)i + k]); *p++ = ' '; } break;
				case K_F3: for (int k = 0; k < 3; k++) { p = put_f6(p, (double)(float)A.d[3 * (size_t)i + k]); *p++ = ' '; } break;
				case K_F1: p = put_f6(p, (double)(float)A.d[i]); *p++ = ' '; break;
				case K_I1: p = put_int(p, A.i[i]); *p++ = ' '; break;
				case K_SEQ0: p = put_int(p, e); *p++ = ' '; break;
				case K_SEQ1: p = put_int(p, e + 1); *p++ = ' '; break;
				case K_ONE: *p++ = '1'; *p++ = ' '; break;
			}
		}
		buf.resize((size_t)(p - &buf[0]));
	}
	bool array(const Arr& A) const {
		fputs(A.open, fp);
		const int m = count(), BLK = 16384;
		const int nBlocks = (m + BLK - 1) / BLK;
		// rounds of nThreads * 8 blocks: formatted in parallel (blocks handed out by a counter), written in order
		const int perRound = nThreads * 8;
		std::vector<std::string> bufs((size_t)std::min(perRound, std::max(nBlocks, 1)));
		for (int b0 = 0; b0 < nBlocks; b0 += perRound) {
			const int nb = std::min(perRound, nBlocks - b0);
			std::atomic<int> next(0);
			auto work = [&]() { for (int k = next.fetch_add(1); k < nb; k = next.fetch_add(1)) { const int a = (b0 + k) * BLK; format(A, a, std::min(m, a + BLK), bufs[(size_t)k]); } };
			std::vector<std::thread> th;
			for (int t = 1; t < std::min(nThreads, nb); t++) th.emplace_back(work);
			work();
			for (auto& t : th) t.join();
			for (int k = 0; k < nb; k++) if (fwrite(bufs[(size_t)k].data(), 1, bufs[(size_t)k].size(), fp) != bufs[(size_t)k].size()) return false;
		}
		fputs("\n        </DataArray>\n", fp);
		return true;
	}
};
}  // namespace

extern "C" int pmps_write_vtu(pmps_t* h, const char* filename, unsigned flags) {
	if (!h || !filename) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_write_vtu: no particle state uploaded");
	const PmpsParams& P = h->K.p;
	const bool fsOnly = flags & PMPS_VTU_FREE_SURFACE_ONLY, rheo = (flags & PMPS_VTU_NONNEWTONIAN) != 0;
	if (rheo && P.fluidType != 1) return pmps_fail(h, PMPS_E_STATE, "pmps_write_vtu: PMPS_VTU_NONNEWTONIAN on a handle created with fluidType 0 (the reference allocates those arrays only for fluidType 1)");
	const size_t n0 = (size_t)h->numParticlesZero;
	// only what the file holds comes back (host order)
	std::vector<double> pos(3 * n0), vel(3 * n0), press(n0), pndi, pndSmall, dev, pnorm, conc, gconc, RHO, Cv, MEU, MEUY, II, psm, inert, prn, VF;
	std::vector<int32_t> bc(n0), ptype, numNeigh, PTYPE, nearMesh;
	PmpsHostArrays a; memset(&a, 0, sizeof a);
	a.numParticlesZero = (int)n0;
	a.pos = pos.data(); a.vel = vel.data(); a.press = press.data(); a.particleBC = bc.data();
	if (fsOnly || (flags & PMPS_VTU_AUXILIAR)) { ptype.resize(n0); a.particleType = ptype.data(); }
	if (flags & PMPS_VTU_PND) { pndi.resize(n0); pndSmall.resize(n0); a.pndi = pndi.data(); a.pndSmall = pndSmall.data(); }
	if (flags & PMPS_VTU_NEIGH) { numNeigh.resize(n0); a.numNeigh = numNeigh.data(); }
	if (flags & PMPS_VTU_DEVIATION) { dev.resize(3 * n0); pnorm.resize(3 * n0); a.npcdDeviation = dev.data(); a.polygonNormal = pnorm.data(); }
	if (flags & PMPS_VTU_CONCENTRATION) { conc.resize(n0); gconc.resize(3 * n0); a.concentration = conc.data(); a.gradConcentration = gconc.data(); }
	if (rheo) {
		RHO.resize(n0); Cv.resize(n0); MEU.resize(n0); MEUY.resize(n0); II.resize(n0); psm.resize(n0); PTYPE.resize(n0);
		a.RHO = RHO.data(); a.Cv = Cv.data(); a.MEU = MEU.data(); a.MEU_Y = MEUY.data(); a.II = II.data(); a.p_smooth = psm.data(); a.PTYPE = PTYPE.data();
		if (!fsOnly) { inert.resize(n0); prn.resize(n0); VF.resize(n0); a.Inertia = inert.data(); a.p_rheo_new = prn.data(); a.VF = VF.data(); }
	}
	if ((flags & PMPS_VTU_AUXILIAR) && P.wallType == 1) { nearMesh.resize(n0); a.nearMeshType = nearMesh.data(); }
	{ const int rc = pmps_download(h, &a); if (rc) return rc; }
	const int n = a.numParticles;
	std::vector<int32_t> sel;
	if (fsOnly) { for (int i = 0; i < n; i++) if (bc[i] == P.surface || ptype[i] == P.wall) sel.push_back(i); }   // src/MpsInputOutput.cpp:1598-1602

	std::vector<char> iobuf(8u << 20);   // (outlives the fclose below)
	FILE* fp = fopen(filename, "w");
	if (!fp) return pmps_fail(h, PMPS_E_ARG, "pmps_write_vtu: cannot open %s", filename);
	setvbuf(fp, iobuf.data(), _IOFBF, iobuf.size());
	Writer W; W.fp = fp; W.n = n; W.sel = fsOnly ? &sel : nullptr;
	W.nThreads = (int)std::thread::hardware_concurrency(); if (W.nThreads < 1) W.nThreads = 1; if (W.nThreads > 64) W.nThreads = 64;
	if (const char* e = getenv("PMPS_VTU_THREADS")) { const int t = atoi(e); if (t >= 1) W.nThreads = t; }
	const int m = W.count();
	fprintf(fp, "<?xml version='1.0' encoding='UTF-8'?>\n<VTKFile xmlns='VTK' byte_order='LittleEndian' version='0.1' type='UnstructuredGrid'>\n  <UnstructuredGrid>\n");
	fprintf(fp, "    <Piece NumberOfCells='%d' NumberOfPoints='%d'>\n      <Points>\n", m, m);
	bool ok = true;
#define F1A(name, v) ok = ok && W.array(Arr{ "        <DataArray type='Float32' Name='" name "' format='ascii'>\n", K_F1, (v).data(), nullptr })
#define F3A(name, v) ok = ok && W.array(Arr{ "        <DataArray type='Float32' Name='" name "' NumberOfComponents='3' format='ascii'>\n", K_F3, (v).data(), nullptr })
#define I1A(name, v) ok = ok && W.array(Arr{ "        <DataArray type='Int32' Name='" name "' format='ascii'>\n", K_I1, nullptr, (v).data() })
	ok = ok && W.array(Arr{ "        <DataArray type='Float32' Name='Position' NumberOfComponents='3' format='ascii'>\n          ", K_POS3, pos.data(), nullptr });
	fputs("      </Points>\n      <PointData>\n", fp);
	F3A("Velocity", vel); F1A("pressure", press); I1A("BC", bc);
	if (flags & PMPS_VTU_PND) { F1A("pnd", pndi); F1A("pndSmall", pndSmall); }
	if (flags & PMPS_VTU_NEIGH) I1A("nNeigh", numNeigh);
	if (flags & PMPS_VTU_DEVIATION) { F3A("npcdDeviation", dev); F3A("polygonNormal", pnorm); }
	if (flags & PMPS_VTU_CONCENTRATION) { F1A("concentration", conc); F3A("GradCi", gconc); }
	if (rheo) {
		F1A("RHO", RHO); F1A("ConcVol", Cv); F1A("MEU", MEU); F1A("MEUy", MEUY); F1A("II", II); I1A("PTYPE", PTYPE); F1A("p_smooth", psm);
		if (!fsOnly) { F1A("Inertia", inert); F1A("Normal_Stress", prn); F1A("VF", VF); }   // (the free-surface variant stops at p_smooth, :1735-1776)
	}
	if (flags & PMPS_VTU_AUXILIAR) {
		if (!fsOnly) I1A("ParticleType", ptype);                                            // (and has no ParticleType array, :1777-1790)
		if (P.wallType == 1) I1A("nearMeshType", nearMesh);
	}
	fputs("      </PointData>\n      <Cells>\n", fp);
	ok = ok && W.array(Arr{ "        <DataArray type='Int32' Name='connectivity' format='ascii'>\n", K_SEQ0, nullptr, nullptr });
	ok = ok && W.array(Arr{ "        <DataArray type='Int32' Name='offsets' format='ascii'>\n", K_SEQ1, nullptr, nullptr });
	ok = ok && W.array(Arr{ "        <DataArray type='UInt8' Name='types' format='ascii'>\n", K_ONE, nullptr, nullptr });
#undef F1A
#undef F3A
#undef I1A
	fputs("      </Cells>\n    </Piece>\n  </UnstructuredGrid>\n</VTKFile>\n", fp);
	ok = ok && !ferror(fp);
	if (fclose(fp) != 0) ok = false;
	if (!ok) return pmps_fail(h, PMPS_E_ARG, "pmps_write_vtu: write to %s failed", filename);
	return PMPS_OK;
}
