/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the voxel-model and dynamic-light passes of the software renderer.
 * Used by tests/ (and nothing in the product).  Pinned: tests/test_oracle_models.py compares it bit-for-bit with the
 * unmodified reference compiled in place (oracle/_ref) and with tests/golden/models_title.npz.
 *
 * Follows, in the reference tree /root/reference/Sources:
 *   Draw/SWModel.cpp:25-100          SWModel::SWModel            -> orc_model_create
 *   Draw/SWModelRenderer.cpp:46-277  SWModelRenderer::RenderInner -> render_model
 *   Client/SceneDefinition.cpp:26-72 projection / view matrices   -> scene_matrices
 *   Core/Math.cpp:108-181            Matrix4 * Matrix4, Matrix4 * Vector4
 *   Draw/SWRenderer.cpp:207-227      BuildFrustrum,  :1130-1147 SphereFrustrumCull
 *   Draw/SWRenderer.cpp:578-649      AddLight                     -> add_light
 *   Draw/SWRenderer.cpp:229-339      ApplyDynamicLight            -> apply_light
 *   Draw/SWRenderer.cpp:709-745      EndScene order               -> orc_render_scene
 */
#include "osb_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>

int orc_render_terrain(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h, int numerics,
                       int threads, int under, uint32_t *bgrx, float *depth);
void orc_apply_fog(const orc_scene *scene, const float fogColor[3], float fogDist, int numerics, int threads, int w, int h,
                   uint32_t *bgrx, float *depth);

struct orc_model {
	int w, h, d;
	float origin[3];
	float center[3], radius; /* SWModel.cpp:28-32 */
	uint32_t *data, *addr;
	size_t nData, nAddr;
};

static void push(uint32_t **a, size_t *n, size_t *cap, uint32_t v) {
	if (*n == *cap) {
		*cap = *cap ? *cap * 2 : 1024;
		*a = (uint32_t *)realloc(*a, *cap * sizeof(uint32_t));
	}
	(*a)[(*n)++] = v;
}

/* VoxelModel::IsSolid, Core/VoxelModel.h:126-132 */
static int vm_solid(const uint64_t *solid, int w, int h, int d, int x, int y, int z) {
	if (x < 0 || y < 0 || x >= w || y >= h || z < 0 || z >= d)
		return 0;
	return (int)((solid[x + y * w] >> z) & 1);
}

/* SWModel.cpp:25-100.  solid[x + y*w] bit z, colors[(x + y*w)*d + z] as VoxelModel stores them. */
orc_model *orc_model_create(int w, int h, int d, const float origin[3], const uint64_t *solid, const uint32_t *colors) {
	if (w <= 0 || h <= 0 || d <= 0 || d > 64)
		return NULL;
	orc_model *m = (orc_model *)calloc(1, sizeof(*m));
	m->w = w;
	m->h = h;
	m->d = d;
	memcpy(m->origin, origin, sizeof(m->origin));
	m->center[0] = (float)w;
	m->center[1] = (float)h;
	m->center[2] = (float)d;
	for (int k = 0; k < 3; k++)
		m->center[k] *= 0.5f;
	m->radius = sqrtf(m->center[0] * m->center[0] + m->center[1] * m->center[1] + m->center[2] * m->center[2]);
	size_t capD = 0, capA = 0;
	for (int y = 0; y < h; y++) {
		for (int x = 0; x < w; x++) {
			push(&m->addr, &m->nAddr, &capA, (uint32_t)m->nData);
			uint64_t map = solid[x + y * w];
			uint64_t map1 = x > 0 ? solid[x - 1 + y * w] : 0;
			uint64_t map2 = x < (w - 1) ? solid[x + 1 + y * w] : 0;
			uint64_t map3 = y > 0 ? solid[x + (y - 1) * w] : 0;
			uint64_t map4 = y < (h - 1) ? solid[x + (y + 1) * w] : 0;
			map1 &= map2;
			map1 &= map3;
			map1 &= map4;
			for (int z = 0; z < d; z++) {
				if (!(map & (1ULL << z)))
					continue;
				if (z == 0 || z == (d - 1) || ((map >> (z - 1)) & 7ULL) != 7ULL || (map1 & (1ULL << z)) == 0) {
					uint32_t col = colors[(size_t)(x + y * w) * d + z];
					uint32_t enc = (col & 0xff00) | ((col & 0xff) << 16) | ((col & 0xff0000) >> 16);
					enc |= (uint32_t)z << 24;
					push(&m->data, &m->nData, &capD, enc);
					uint32_t normal;
					if ((col >> 24) == 1) { /* MaterialType::Emissive, Core/VoxelModel.h */
						normal = 27;
					} else {
						int nx = 0, ny = 0, nz = 0;
						for (int cx = -1; cx <= 1; cx++)
							for (int cy = -1; cy <= 1; cy++)
								for (int cz = -1; cz <= 1; cz++) {
									if (vm_solid(solid, w, h, d, x + cx, y + cy, z + cz)) {
										nx -= cx;
										ny -= cy;
										nz -= cz;
									} else {
										nx += cx;
										ny += cy;
										nz += cz;
									}
								}
						nx = (nx > 1 ? 1 : (nx < -1 ? -1 : nx)) + 1;
						ny = (ny > 1 ? 1 : (ny < -1 ? -1 : ny)) + 1;
						nz = (nz > 1 ? 1 : (nz < -1 ? -1 : nz)) + 1;
						normal = (uint32_t)(nx + ny * 3 + nz * 9);
					}
					push(&m->data, &m->nData, &capD, normal);
				}
			}
			push(&m->data, &m->nData, &capD, 0xffffffffU);
		}
	}
	return m;
}

void orc_model_free(orc_model *m) {
	if (!m)
		return;
	free(m->data);
	free(m->addr);
	free(m);
}

int orc_model_render_data(const orc_model *m, uint32_t *data, int dataCap, uint32_t *addr, int addrCap) {
	if (!m)
		return -1;
	if (data && dataCap >= (int)m->nData)
		memcpy(data, m->data, m->nData * 4);
	if (addr && addrCap >= (int)m->nAddr)
		memcpy(addr, m->addr, m->nAddr * 4);
	return (int)m->nData;
}

/* ---- numerics of the build (Draw/SWFeatureLevel.h:89-121) ---- */
static inline float m_rcp(int sse, float b) { return sse ? _mm_cvtss_f32(_mm_rcp_ss(_mm_set_ss(b))) : 1.f / b; }
static inline float m_rsqrt(int sse, float b) { return sse ? _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(b))) : 1.f / sqrtf(b); }

static inline float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline float fmax_std(float a, float b) { return a < b ? b : a; } /* std::max */
static inline int to_fixed8(float v) {                                     /* SWUtils.h:68-71 */
	int i = (int)(v * 255.f + .5f);
	return i > 255 ? 255 : (i < 0 ? 0 : i);
}
static inline int to_fixed_factor8(float v) { /* SWUtils.h:73-76 */
	int i = (int)(v * 256.f + .5f);
	return i > 256 ? 256 : (i < 0 ? 0 : i);
}

typedef struct {
	float viewproj[16];  /* column-major like Matrix4::m */
	float frustum[6][4]; /* Plane3 n, w */
} scene_mats;

/* Matrix4 * (x, y, z, w), Core/Math.cpp:162-181 */
static void mat_vec4(const float *m, const float v[3], float w, float out[4]) {
	for (int r = 0; r < 4; r++) {
		float acc = m[r] * v[0];
		acc += m[4 + r] * v[1];
		acc += m[8 + r] * v[2];
		acc += m[12 + r] * w;
		out[r] = acc;
	}
}

static void plane_through(float out[4], const float point[3], const float n[3]) { /* Plane3::PlaneWithPointOnPlane */
	out[0] = n[0];
	out[1] = n[1];
	out[2] = n[2];
	out[3] = -dot3(point, n);
}

static void scene_matrices(const orc_scene *s, scene_mats *S) {
	const float near = s->zNear, far = s->zFar;
	const float t = near * tanf(s->fovY * .5f), r = near * tanf(s->fovX * .5f);
	const float a = r * 2.f, b = t * 2.f, c = far - near;
	float P[16], V[16];
	memset(P, 0, sizeof(P));
	memset(V, 0, sizeof(V));
	P[0] = near * 2.f / a;
	P[5] = near * 2.f / b;
	P[10] = -(far + near) / c;
	P[11] = -1.f;
	P[14] = -(far * near * 2.f) / c;
	V[15] = 1.f;
	for (int k = 0; k < 3; k++) {
		V[k * 4 + 0] = s->axis[0][k];
		V[k * 4 + 1] = s->axis[1][k];
		V[k * 4 + 2] = -s->axis[2][k];
	}
	float v[4];
	mat_vec4(V, s->origin, 1.f, v); /* translation column still zero here */
	V[12] = -v[0];
	V[13] = -v[1];
	V[14] = -v[2];
	/* projectionMatrix * viewMatrix (Math.cpp:108-128) */
	for (int col = 0; col < 4; col++)
		for (int row = 0; row < 4; row++)
			S->viewproj[col * 4 + row] =
			  V[col * 4] * P[row] + V[col * 4 + 1] * P[4 + row] + V[col * 4 + 2] * P[8 + row] + V[col * 4 + 3] * P[12 + row];
	plane_through(S->frustum[0], s->origin, s->axis[2]);
	for (int k = 0; k < 4; k++)
		S->frustum[1][k] = -S->frustum[0][k];
	S->frustum[0][3] -= s->zNear;
	S->frustum[1][3] += s->zFar;
	const float xCos = cosf(s->fovX * .5f), xSin = sinf(s->fovX * .5f);
	const float yCos = cosf(s->fovY * .5f), ySin = sinf(s->fovY * .5f);
	float n[3];
	for (int k = 0; k < 3; k++)
		n[k] = s->axis[2][k] * xSin - s->axis[0][k] * xCos;
	plane_through(S->frustum[2], s->origin, n);
	for (int k = 0; k < 3; k++)
		n[k] = s->axis[2][k] * xSin + s->axis[0][k] * xCos;
	plane_through(S->frustum[3], s->origin, n);
	for (int k = 0; k < 3; k++)
		n[k] = s->axis[2][k] * ySin - s->axis[1][k] * yCos;
	plane_through(S->frustum[4], s->origin, n);
	for (int k = 0; k < 3; k++)
		n[k] = s->axis[2][k] * ySin + s->axis[1][k] * yCos;
	plane_through(S->frustum[5], s->origin, n);
}

/* SWModelRenderer.cpp:46-277 */
static void render_model(const orc_model *model, const float *mat, const float customColorF[3], const orc_scene *sc,
                         const scene_mats *S, int sse, int fw, int fh, uint32_t *fb, float *db) {
	float origin[3] = {mat[12], mat[13], mat[14]};
	const float *axis1 = mat, *axis2 = mat + 4, *axis3 = mat + 8;
	float rmo[3];
	for (int k = 0; k < 3; k++)
		rmo[k] = model->origin[k] + 0.25f;
	for (int k = 0; k < 3; k++)
		origin[k] += axis1[k] * rmo[0];
	for (int k = 0; k < 3; k++)
		origin[k] += axis2[k] * rmo[1];
	for (int k = 0; k < 3; k++)
		origin[k] += axis3[k] * rmo[2];
	const int w = model->w, h = model->h;

	uint8_t brights[28];
	{
		const float lightVec[3] = {0.f, -0.707f, -0.707f};
		float dot1 = dot3(axis1, lightVec) * m_rsqrt(sse, dot3(axis1, axis1));
		float dot2 = dot3(axis2, lightVec) * m_rsqrt(sse, dot3(axis2, axis2));
		float dot3v = dot3(axis3, lightVec) * m_rsqrt(sse, dot3(axis3, axis3));
		for (int x = 0; x < 3; x++) {
			float d = 0.f;
			int cnt = 0;
			if (x == 0) {
				d = -dot1;
				cnt = 1;
			} else if (x == 2) {
				d = dot1;
				cnt = 1;
			}
			for (int y = 0; y < 3; y++) {
				float d2 = d; /* computed and not read again, as in the reference (:96-108) */
				int cnt2 = cnt;
				if (y == 0) {
					d2 -= dot2;
					cnt2++;
				} else if (y == 2) {
					d2 += dot2;
					cnt2++;
				}
				(void)d2;
				for (int z = 0; z < 3; z++) {
					float d3 = d; /* :110: restarts from the x term, and the switch below tests y (:112) */
					int cnt3 = cnt2;
					if (y == 0) {
						d3 -= dot3v;
						cnt3++;
					} else if (y == 2) {
						d3 += dot3v;
						cnt3++;
					}
					if (cnt3 == 2)
						d3 *= 0.707f;
					else if (cnt3 == 3)
						d3 *= 0.57735f;
					d3 = 192.f + d3 * 62.f;
					brights[x + y * 3 + z * 9] = (uint8_t)(int)d3;
				}
			}
		}
	}
	brights[27] = 255;

	{
		float center[3] = {origin[0], origin[1], origin[2]};
		for (int k = 0; k < 3; k++)
			center[k] += axis1[k] * model->center[0];
		for (int k = 0; k < 3; k++)
			center[k] += axis2[k] * model->center[1];
		for (int k = 0; k < 3; k++)
			center[k] += axis3[k] * model->center[2];
		float largestAxis = dot3(axis1, axis1);
		largestAxis = fmax_std(largestAxis, dot3(axis2, axis2));
		largestAxis = fmax_std(largestAxis, dot3(axis3, axis3));
		const float radius = model->radius * sqrtf(largestAxis);
		for (int i = 0; i < 6; i++) /* SphereFrustrumCull */
			if (dot3(S->frustum[i], center) + S->frustum[i][3] < -radius)
				return;
	}

	const float sx = fw * 0.5f, sy = -fh * 0.5f;
	const int offX = fw >> 1, offY = fh >> 1;
	float tOrigin[4], tAxis1[4], tAxis2[4], tAxis3[4];
	mat_vec4(S->viewproj, origin, 1.f, tOrigin);
	mat_vec4(S->viewproj, axis1, 0.f, tAxis1);
	mat_vec4(S->viewproj, axis2, 0.f, tAxis2);
	mat_vec4(S->viewproj, axis3, 0.f, tAxis3);
	float *all[4] = {tOrigin, tAxis1, tAxis2, tAxis3};
	for (int k = 0; k < 4; k++) {
		all[k][0] *= sx;
		all[k][1] *= sy;
		all[k][2] *= 1.f;
		all[k][3] *= 1.f;
	}
	float pointDiameter;
	{
		float l1 = tAxis1[0] * tAxis1[0] + tAxis1[1] * tAxis1[1] + tAxis1[2] * tAxis1[2] + tAxis1[3] * tAxis1[3];
		float l2 = tAxis2[0] * tAxis2[0] + tAxis2[1] * tAxis2[1] + tAxis2[2] * tAxis2[2] + tAxis2[3] * tAxis2[3];
		float l3 = tAxis3[0] * tAxis3[0] + tAxis3[1] * tAxis3[1] + tAxis3[2] * tAxis3[2] + tAxis3[3] * tAxis3[3];
		float largestAxis = l1;
		largestAxis = fmax_std(largestAxis, l2);
		largestAxis = fmax_std(largestAxis, l3);
		pointDiameter = sqrtf(largestAxis);
	}
	const uint32_t customColor = (uint32_t)to_fixed8(customColorF[2]) | ((uint32_t)to_fixed8(customColorF[1]) << 8) |
	                             ((uint32_t)to_fixed8(customColorF[0]) << 16);
	float v1[4] = {tOrigin[0], tOrigin[1], tOrigin[2], tOrigin[3]};
	const float zNear = sc->zNear;
	for (int x = 0; x < w; x++) {
		float v2[4] = {v1[0], v1[1], v1[2], v1[3]};
		for (int y = 0; y < h; y++) {
			const uint32_t *mp = model->data + model->addr[x + y * w];
			while (*mp != 0xffffffffU) {
				uint32_t data = *(mp++);
				uint32_t normal = *(mp++);
				int z = (int)(data >> 24);
				float vv[4];
				for (int k = 0; k < 4; k++)
					vv[k] = v2[k] + tAxis3[k] * (float)z;
				if (vv[2] < zNear)
					continue;
				float zval = vv[2];
				vv[2] = pointDiameter;
				float scl = m_rcp(sse, vv[3]);
				for (int k = 0; k < 4; k++)
					vv[k] *= scl;
				int ix = (int)vv[0] + offX;
				int iy = (int)vv[1] + offY;
				int idm = (int)(vv[2] + .99f);
				idm = idm > 1 ? idm : 1;
				int minX = ix - (idm >> 1);
				int minY = iy - (idm >> 1);
				if (minX >= fw || minY >= fh)
					continue;
				int maxX = (int)((uint32_t)ix + (uint32_t)idm); /* wraps like the compiled int add */
				int maxY = (int)((uint32_t)iy + (uint32_t)idm);
				if (maxX <= 0 || maxY <= 0)
					continue;
				minX = minX > 0 ? minX : 0;
				minY = minY > 0 ? minY : 0;
				maxX = maxX < fw ? maxX : fw;
				maxY = maxY < fh ? maxY : fh;
				uint32_t color = data & 0xffffff;
				if (color == 0)
					color = customColor;
				int bright = brights[normal < 28 ? normal : 27];
				/* the SSE2 branch (mulhi_epu16 by bright << 8) and the scalar branch give the same bytes */
				uint32_t c1 = color & 0xff00;
				uint32_t c2 = color & 0xff00ff;
				c1 *= (uint32_t)bright;
				c2 *= (uint32_t)bright;
				color = ((c1 & 0xff0000) | (c2 & 0xff00ff00)) >> 8;
				for (int yy = minY; yy < maxY; yy++)
					for (int xx = minX; xx < maxX; xx++) {
						size_t o = (size_t)yy * fw + xx;
						if (zval < db[o]) {
							db[o] = zval;
							fb[o] = color;
						}
					}
			}
			for (int k = 0; k < 4; k++)
				v2[k] += tAxis2[k];
		}
		for (int k = 0; k < 4; k++)
			v1[k] += tAxis1[k];
	}
}

typedef struct {
	int minX, minY, maxX, maxY;
	orc_light p;
} dyn_light;

static int view_trans(float v, float fov, float screenSize) {
	v = (v / fov) * 0.5f + 0.5f;
	v = v < 0.f ? 0.f : v; /* std::max(v, 0.f) */
	v = 1.f < v ? 1.f : v; /* std::min(v, 1.f) */
	v *= screenSize;
	return (int)v;
}
static void view_range(float cx, float cy, float fov, float screenSize, int out[2]) {
	float dist = cx * cx + cy * cy - 1.f;
	if (dist <= 0.f) {
		out[0] = 0;
		out[1] = (int)screenSize;
		return;
	}
	dist = sqrtf(dist);
	if (cx <= 1.f) {
		if (cy > 0.f) {
			out[0] = view_trans(cx * cy - dist, fov, screenSize);
			out[1] = (int)screenSize;
		} else {
			out[0] = 0;
			out[1] = view_trans(cx * cy + dist, fov, screenSize);
		}
	} else {
		out[0] = view_trans(cx * cy - dist, fov, screenSize);
		out[1] = view_trans(cx * cy + dist, fov, screenSize);
	}
}

/* SWRenderer.cpp:578-649; returns 0 when the light is culled */
static int add_light(const orc_scene *s, const orc_light *p, int fw, int fh, float fogDistance, dyn_light *L) {
	float diff[3];
	for (int k = 0; k < 3; k++)
		diff[k] = p->origin[k] - s->origin[k];
	float rad2 = p->radius * p->radius;
	float poweredLength = dot3(diff, diff);
	float fogCullRange = p->radius + fogDistance;
	if (poweredLength > fogCullRange * fogCullRange)
		return 0;
	if (poweredLength < rad2) {
		L->minX = 0;
		L->minY = 0;
		L->maxX = fw;
		L->maxY = fh;
	} else if (dot3(diff, s->axis[2]) < 0.f) {
		return 0;
	} else {
		float invRad = 1.f / p->radius;
		int rx[2], ry[2];
		/* both ranges take viewAxis[0] as the second coordinate (:637-642) */
		view_range(dot3(diff, s->axis[2]) * invRad, dot3(diff, s->axis[0]) * invRad, tanf(s->fovX * 0.5f), (float)fw, rx);
		view_range(dot3(diff, s->axis[1]) * invRad, dot3(diff, s->axis[0]) * invRad, tanf(s->fovY * 0.5f), (float)fh, ry);
		L->minX = rx[0];
		L->maxX = rx[1];
		L->minY = ry[0];
		L->maxY = ry[1];
	}
	L->p = *p;
	return 1;
}

/* SWRenderer.cpp:229-339, worker threads emulated one after the other (their row bands are disjoint) */
static void apply_light(const orc_scene *s, const dyn_light *L, int threads, int fw, int fh, uint32_t *fbAll, float *dbAll) {
	const float fovX = tanf(s->fovX * 0.5f), fovY = tanf(s->fovY * 0.5f);
	const float dvx = -fovX * 2.f / (float)fw, dvy = -fovY * 2.f / (float)fh;
	const int minX = L->minX, minY = L->minY, maxX = L->maxX, maxY = L->maxY;
	const int lightHeight = maxY - minY;
	float diff[3], lightCenter[3];
	for (int k = 0; k < 3; k++)
		diff[k] = L->p.origin[k] - s->origin[k];
	for (int k = 0; k < 3; k++)
		lightCenter[k] = dot3(diff, s->axis[k]);
	const int lightR = to_fixed_factor8(L->p.color[0]), lightG = to_fixed_factor8(L->p.color[1]),
	          lightB = to_fixed_factor8(L->p.color[2]);
	const float invRadius2 = 1.f / (L->p.radius * L->p.radius);
	for (unsigned threadId = 0; threadId < (unsigned)threads; threadId++) {
		/* int * unsigned -> unsigned arithmetic, as written in the reference lambda (:262-263) */
		int startY = (int)((unsigned)lightHeight * threadId / (unsigned)threads);
		int endY = (int)((unsigned)lightHeight * (threadId + 1) / (unsigned)threads);
		startY += minY;
		endY += minY;
		uint32_t *fb = fbAll + (ptrdiff_t)startY * fw + minX;
		float *db = dbAll + (ptrdiff_t)startY * fw + minX;
		float vy = fovY + dvy * startY;
		float vx = fovX + dvx * minX;
		const int lightWidth = maxX - minX;
		for (int y = startY; y < endY; y++) {
			float vx2 = vx;
			uint32_t *fb2 = fb;
			float *db2 = db;
			for (int x = lightWidth; x > 0; x--) {
				float pos[3];
				pos[2] = *db2;
				pos[0] = vx2 * pos[2];
				pos[1] = vy * pos[2];
				for (int k = 0; k < 3; k++)
					pos[k] -= lightCenter[k];
				float dist = dot3(pos, pos);
				dist *= invRadius2;
				if (dist < 1.f) {
					float strength = 1.f - dist;
					strength *= strength;
					strength *= 256.f;
					int factor = (int)strength;
					int aR = lightR * factor, aG = lightG * factor, aB = lightB * factor;
					uint32_t src = *fb2;
					uint32_t sR = (src >> 16) & 0xff, sG = (src >> 8) & 0xff, sB = src & 0xff;
					uint32_t pR = (uint32_t)aR * sR, pG = (uint32_t)aG * sG, pB = (uint32_t)aB * sB; /* int *= unsigned */
					uint32_t dR = (uint32_t)((int)pR >> 16), dG = (uint32_t)((int)pG >> 16), dB = (uint32_t)((int)pB >> 16);
					dR = dR + sR < 255 ? dR + sR : 255;
					dG = dG + sG < 255 ? dG + sG : 255;
					dB = dB + sB < 255 ? dB + sB : 255;
					*fb2 = dB | (dG << 8) | (dR << 16);
				}
				vx2 += dvx;
				fb2++;
				db2++;
			}
			vy += dvy;
			fb += fw;
			db += fw;
		}
	}
}

/* SWRenderer::EndScene (SWRenderer.cpp:709-745) without sprites: clear, terrain, models in submission order, lights in
 * submission order, fog.  solid may be NULL (no map set: the terrain pass is skipped, :720). */
int orc_render_scene(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h,
                     const float fogColor[3], float fogDist, int numerics, int threads, int under,
                     const orc_model_instance *inst, int nInst, const orc_light *lights, int nLights, uint32_t *bgrx,
                     float *depth) {
	if ((w & 7) || (h & 7) || w <= 0 || h <= 0)
		return -1;
	if (threads < 1)
		threads = 1;
	if (threads > 32)
		threads = 32;
	if (solid) {
		int rc = orc_render_terrain(solid, color, scene, w, h, numerics, threads, under, bgrx, depth);
		if (rc)
			return rc;
	} else {
		for (size_t i = 0; i < (size_t)w * h; i++)
			bgrx[i] = 0x7f7f7f;
	}
	const int sse = numerics != ORC_NUM_IEEE;
	scene_mats S;
	scene_matrices(scene, &S);
	/* AddLight runs at submission time, before EndScene; it reads only the scene and the fog distance */
	dyn_light *kept = (dyn_light *)malloc(sizeof(dyn_light) * (size_t)(nLights > 0 ? nLights : 1));
	int nKept = 0;
	for (int i = 0; i < nLights; i++)
		if (add_light(scene, &lights[i], w, h, fogDist, &kept[nKept]))
			nKept++;
	for (int i = 0; i < nInst; i++)
		render_model(inst[i].model, inst[i].matrix, inst[i].customColor, scene, &S, sse, w, h, bgrx, depth);
	for (int i = 0; i < nKept; i++)
		apply_light(scene, &kept[i], threads, w, h, bgrx, depth);
	free(kept);
	orc_apply_fog(scene, fogColor, fogDist, numerics, threads, w, h, bgrx, depth);
	return 0;
}
