// C wrappers over the C++ host layer (VoxelMap, B200Renderer) so that the Python tests and bench.py
// can drive the same objects a C++ integrator would use.  Not part of the osb200 C ABI proper.
#include <cstring>
#include <exception>
#include <string>

#include "../../include/osb200.h"
#include "VoxelMap.h"

#define OSBH_API extern "C" __attribute__((visibility("default")))

namespace {
	thread_local std::string g_err;
	template <class F> int guarded(F f) {
		try {
			f();
			return 0;
		} catch (const std::exception &e) {
			g_err = e.what();
			return -1;
		}
	}
} // namespace

OSBH_API const char *osbh_last_error() { return g_err.c_str(); }

OSBH_API void *osbh_map_new() { return new osb200::VoxelMap(); }
OSBH_API void *osbh_map_load(const uint8_t *data, size_t len) {
	osb200::VoxelMap *m = nullptr;
	guarded([&] { m = osb200::VoxelMap::Load(data, len).release(); });
	return m;
}
OSBH_API void *osbh_map_load_file(const char *path) {
	osb200::VoxelMap *m = nullptr;
	guarded([&] { m = osb200::VoxelMap::LoadFile(path).release(); });
	return m;
}
OSBH_API void osbh_map_free(void *m) { delete static_cast<osb200::VoxelMap *>(m); }
OSBH_API const uint64_t *osbh_map_solid(void *m) { return static_cast<osb200::VoxelMap *>(m)->SolidData(); }
OSBH_API const uint32_t *osbh_map_color(void *m) { return static_cast<osb200::VoxelMap *>(m)->ColorData(); }
OSBH_API int osbh_map_set(void *m, int x, int y, int z, int solid, uint32_t color) {
	return guarded([&] { static_cast<osb200::VoxelMap *>(m)->Set(x, y, z, solid != 0, color); });
}
OSBH_API int osbh_map_set_arrays(void *m, const uint64_t *solid, const uint32_t *color) {
	auto *map = static_cast<osb200::VoxelMap *>(m);
	std::memcpy(map->MutableSolidData(), solid, sizeof(uint64_t) * 512 * 512);
	std::memcpy(map->MutableColorData(), color, sizeof(uint32_t) * 512 * 512 * 64);
	return 0;
}
