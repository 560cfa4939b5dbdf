// Brute-force check of openspades_b200/csrc/libm_f32.cuh against the box's libm (glibc).
// Prints "<name> <mismatches> <checked>" per function; exit code 0 iff all mismatch counts are 0.
#include <cmath>
#include <cstdint>
#include <cstdio>

#include "../../openspades_b200/csrc/libm_f32.cuh"

int main() {
	long badAtan = 0, nAtan = 0, badTan = 0, nTan = 0;
#pragma omp parallel for reduction(+ : badAtan, nAtan, badTan, nTan) schedule(static, 1 << 16)
	for (int64_t i = 0; i < (int64_t)0x7f800001; i++) {
		const uint32_t u = (uint32_t)i;
		for (uint32_t sign = 0; sign <= 0x80000000u; sign += 0x80000000u) {
			const float x = osbm::u2f(u | sign);
			const float a = atanf(x), b = osbm::atanf_glibc(x);
			nAtan++;
			if (osbm::f2u(a) != osbm::f2u(b) && !(a != a && b != b))
				badAtan++;
			if (u < 0x3fcccccdu) { // |x| < 1.6 covers SpecialTan's (-pi/2, pi/2)
				const float t = tanf(x), t2 = osbm::tanf_glibc(x);
				nTan++;
				if (osbm::f2u(t) != osbm::f2u(t2))
					badTan++;
			}
			if (sign)
				break;
		}
	}
	printf("atanf %ld %ld\ntanf %ld %ld\n", badAtan, nAtan, badTan, nTan);
	return (badAtan || badTan) ? 1 : 0;
}
