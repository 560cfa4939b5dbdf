// Host-side mirror of client::GameMapWrapper (Sources/Client/GameMapWrapper.h:71-117): floating-block detection for
// a VoxelMap whose device copy lives in an osb200 context.  Same three operations -- Rebuild, AddBlock, RemoveBlocks --
// with the connectivity itself computed on the GPU (osb200_links_*); the class applies the edits to the host map
// (VoxelMap::Set, so listeners fire as with GameMap::Set), flushes the dirty columns and asks the device.
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/osb200.h"
#include "VoxelMap.h"

namespace osb200 {
	// client::CellPos (GameMapWrapper.h:30-49)
	struct CellPos {
		short x, y, z;
		CellPos() = default;
		CellPos(int xx, int yy, int zz) : x((short)xx), y((short)yy), z((short)zz) {}
		bool operator<(const CellPos &p) const { return x != p.x ? x < p.x : (y != p.y ? y < p.y : z < p.z); }
		bool operator==(const CellPos &p) const { return x == p.x && y == p.y && z == p.z; }
	};

	class VoxelMapWrapper {
	public:
		// ctx must hold `map` (osb200_map_upload / B200Renderer::SetGameMap); neither is owned
		VoxelMapWrapper(VoxelMap &map, osb200_ctx *ctx) : map_(map), ctx_(ctx) {}

		// GameMapWrapper::Rebuild (GameMapWrapper.cpp:50-102)
		void Rebuild();
		// GameMapWrapper::AddBlock (:104-193): adds the voxel and links it if it touches the linked set
		void AddBlock(int x, int y, int z, uint32_t color);
		// the same for many voxels with one device round trip
		void AddBlocks(const std::vector<CellPos> &cells, const std::vector<uint32_t> &colors);
		// GameMapWrapper::RemoveBlocks (:199-361): removes the cells, returns the blocks that now float (sorted by
		// CellPos::operator<; the reference returns the same set in walk order).  It does not remove them.
		std::vector<CellPos> RemoveBlocks(const std::vector<CellPos> &cells);

		// World::ClusterizeBlocks (World.cpp:154-226), the next step of World::ApplyBlockActions (:249): the floating
		// set split into its 6-connected pieces, one falling object each.  Pieces come ordered by their smallest cell,
		// cells inside a piece sorted (the reference: unordered_map iteration order / walk order; the same partition).
		std::vector<std::vector<CellPos>> ClusterizeBlocks(const std::vector<CellPos> &blocks);

	private:
		void Flush();
		VoxelMap &map_;
		osb200_ctx *ctx_;
	};
} // namespace osb200
