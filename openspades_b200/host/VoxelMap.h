// Host-side voxel map with client::GameMap's surface (Sources/Client/GameMap.h:38-167): same storage
// order, same Set()/listener contract, same VXL decoder.  It is the host truth that feeds
// osb200_map_upload / osb200_map_flush; ray queries and rendering run on the GPU through the C ABI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <list>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <vector>

struct osb200_ctx;
struct osb200_rayhit;

namespace osb200 {

	class VoxelMap;

	// client::IGameMapListener (Sources/Client/IGameMapListener.h:26-29)
	class IVoxelMapListener {
	public:
		virtual ~IVoxelMapListener() {}
		virtual void GameMapChanged(int x, int y, int z, VoxelMap *) = 0;
	};

	class VoxelMap {
	public:
		enum { DefaultWidth = 512, DefaultHeight = 512, DefaultDepth = 64 };

		VoxelMap(); // flat ground at z = 63 like GameMap::GameMap (GameMap.cpp:36-49), deterministic colour
		// GameMap::Load (GameMap.cpp:519-596): VOXLAP5 span stream -> map. Throws std::runtime_error on truncation.
		static std::unique_ptr<VoxelMap> Load(const uint8_t *data, size_t len);
		static std::unique_ptr<VoxelMap> LoadFile(const char *path);

		int Width() const { return DefaultWidth; }
		int Height() const { return DefaultHeight; }
		int Depth() const { return DefaultDepth; }

		bool IsSolid(int x, int y, int z) const { return ((solid_[Idx(x, y)] >> z) & 1ull) != 0; }
		uint32_t GetColor(int x, int y, int z) const { return color_[Idx(x, y) * 64 + z]; }
		uint64_t GetSolidMapWrapped(int x, int y) const { return solid_[Idx(x & 511, y & 511)]; }
		bool IsSolidWrapped(int x, int y, int z) const { // GameMap.h:91-98
			if (z < 0)
				return false;
			if (z >= DefaultDepth)
				return true;
			return ((solid_[Idx(x & 511, y & 511)] >> z) & 1ull) != 0;
		}
		uint32_t GetColorWrapped(int x, int y, int z) const { return color_[Idx(x & 511, y & 511) * 64 + (z & 63)]; }

		// GameMap::Set (GameMap.h:104-135): notifies listeners synchronously when something changed.
		void Set(int x, int y, int z, bool solid, uint32_t color, bool unsafe = false);

		void AddListener(IVoxelMapListener *);
		void RemoveListener(IVoxelMapListener *);

		// raw arrays in the reference's order: solid[x][y], color[x][y][z]
		const uint64_t *SolidData() const { return solid_.data(); }
		const uint32_t *ColorData() const { return color_.data(); }
		uint64_t *MutableSolidData() { return solid_.data(); }
		uint32_t *MutableColorData() { return color_.data(); }

		// Batched GameMap::CastRay2 (GameMap.cpp:361-507) on the GPU context that holds this map.
		// Throws std::runtime_error on C-ABI failure.
		void CastRay2Batch(osb200_ctx *ctx, const float *v0, const float *dir, int maxSteps,
		                   osb200_rayhit *out, size_t n) const;
		// Batched GameMap::CastRay (GameMap.cpp:208-359): out[i] = {hit, hitBlock.x, .y, .z}
		void CastRayBatch(osb200_ctx *ctx, const float *v0, const float *dir, float length, int32_t *out,
		                  size_t n) const;

	private:
		static size_t Idx(int x, int y) { return (size_t)x * DefaultHeight + (size_t)y; }
		std::vector<uint64_t> solid_;
		std::vector<uint32_t> color_;
		std::list<IVoxelMapListener *> listeners_;
		std::mutex listenersMutex_;
	};
} // namespace osb200
