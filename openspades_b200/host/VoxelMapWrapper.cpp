#include "VoxelMapWrapper.h"

#include <algorithm>
#include <stdexcept>

namespace osb200 {
	static void Check(int rc) {
		if (rc < 0)
			throw std::runtime_error(osb200_last_error());
	}

	void VoxelMapWrapper::Flush() { Check(osb200_map_flush(ctx_, map_.SolidData(), map_.ColorData())); }

	void VoxelMapWrapper::Rebuild() {
		Flush();
		Check(osb200_links_rebuild(ctx_));
	}

	void VoxelMapWrapper::AddBlock(int x, int y, int z, uint32_t color) {
		AddBlocks({CellPos(x, y, z)}, {color});
	}

	void VoxelMapWrapper::AddBlocks(const std::vector<CellPos> &cells, const std::vector<uint32_t> &colors) {
		if (cells.size() != colors.size())
			throw std::invalid_argument("VoxelMapWrapper::AddBlocks: one colour per cell");
		bool any = false;
		for (size_t i = 0; i < cells.size(); i++) {
			const CellPos &p = cells[i];
			if (map_.IsSolid(p.x, p.y, p.z))
				continue; // an existing voxel is already linked or part of a floating cluster (:109-112)
			map_.Set(p.x, p.y, p.z, true, colors[i]);
			Check(osb200_map_mark_dirty(ctx_, p.x, p.y, p.z)); // also when no renderer listens to the map
			any = true;
		}
		if (!any)
			return;
		Flush();
		Check(osb200_links_blocks_added(ctx_));
	}

	std::vector<CellPos> VoxelMapWrapper::RemoveBlocks(const std::vector<CellPos> &cells) {
		if (cells.empty())
			return {};
		std::vector<int16_t> flat;
		flat.reserve(cells.size() * 3);
		for (const CellPos &p : cells) {
			map_.Set(p.x, p.y, p.z, false, 0);
			Check(osb200_map_mark_dirty(ctx_, p.x, p.y, p.z));
			flat.push_back(p.x);
			flat.push_back(p.y);
			flat.push_back(p.z);
		}
		Flush();
		std::vector<int16_t> out(3 * 4096);
		int n = osb200_links_remove_blocks(ctx_, flat.data(), (int)cells.size(), out.data(), (int)(out.size() / 3));
		Check(n);
		if ((size_t)n * 3 > out.size()) { // a large structure fell: fetch the whole result again
			out.resize((size_t)n * 3);
			Check(osb200_links_last_floating(ctx_, out.data(), n));
		}
		std::vector<CellPos> res;
		res.reserve((size_t)n);
		for (int i = 0; i < n; i++)
			res.emplace_back(out[i * 3], out[i * 3 + 1], out[i * 3 + 2]);
		return res;
	}
	std::vector<std::vector<CellPos>> VoxelMapWrapper::ClusterizeBlocks(const std::vector<CellPos> &blocks) {
		if (blocks.empty())
			return {};
		std::vector<int16_t> flat;
		flat.reserve(blocks.size() * 3);
		for (const CellPos &p : blocks) {
			flat.push_back(p.x);
			flat.push_back(p.y);
			flat.push_back(p.z);
		}
		std::vector<int32_t> id(blocks.size());
		const int k = osb200_clusterize_blocks(ctx_, flat.data(), (int)blocks.size(), id.data());
		Check(k);
		std::vector<std::vector<CellPos>> ret((size_t)k);
		for (size_t i = 0; i < blocks.size(); i++)
			ret[(size_t)id[i]].push_back(blocks[i]);
		for (auto &piece : ret) {
			std::sort(piece.begin(), piece.end());
			piece.erase(std::unique(piece.begin(), piece.end()), piece.end()); // a cell named twice is one cell
		}
		return ret;
	}
} // namespace osb200
