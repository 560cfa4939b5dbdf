#include "VoxelMap.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/osb200.h"

namespace osb200 {

	VoxelMap::VoxelMap()
	    : solid_((size_t)DefaultWidth * DefaultHeight, 1ull << 63),
	      color_((size_t)DefaultWidth * DefaultHeight * DefaultDepth, 0x64284067u) {}

	namespace {
		// Cursor over the VXL byte stream with bounds checks (the reference reads through
		// RandomAccessAdaptor, which throws at end of stream).
		struct Reader {
			const uint8_t *p;
			size_t len;
			uint8_t u8(size_t pos) const {
				if (pos >= len)
					throw std::runtime_error("VXL stream truncated");
				return p[pos];
			}
			int s8(size_t pos) const { return (int8_t)u8(pos); }
			uint32_t u32(size_t pos) const {
				if (pos + 3 >= len)
					throw std::runtime_error("VXL stream truncated");
				uint32_t v;
				std::memcpy(&v, p + pos, 4);
				return v;
			}
		};
		// file colour is B,G,R,A bytes; map colour is 0xHHBBGGRR with health 100 (GameMap.cpp:509-517)
		inline uint32_t FileColorToMap(uint32_t c) {
			return ((c >> 16) & 0xff) | (c & 0xff00) | ((c & 0xff) << 16) | (100u << 24);
		}
	} // namespace

	std::unique_ptr<VoxelMap> VoxelMap::Load(const uint8_t *data, size_t len) {
		std::unique_ptr<VoxelMap> map(new VoxelMap());
		std::fill(map->color_.begin(), map->color_.end(), 0u);
		Reader rd{data, len};
		size_t pos = 0;
		// columns are stored y-major in the file
		for (int y = 0; y < DefaultHeight; y++) {
			for (int x = 0; x < DefaultWidth; x++) {
				uint64_t bits = ~0ull; // everything solid until a span says otherwise
				uint32_t *col = &map->color_[Idx(x, y) * 64];
				int airFrom = 0;
				// VXL maps arrive over the network: every z the stream names is range-checked (the reference only
				// SPAsserts, GameMap.h:104-110; the device decoder k_vxl_decode rejects the same streams)
				auto put = [&](int z, uint32_t c) {
					if (z < 0 || z > 63)
						throw std::runtime_error("VXL stream malformed: voxel z outside 0..63");
					col[z] = c;
					bits |= 1ull << z;
				};
				for (int span = 0;; span++) {
					if (span >= 64)
						throw std::runtime_error("VXL stream malformed: more than 64 spans in one column");
					const int chunks = rd.s8(pos);
					const int topStart = rd.s8(pos + 1);
					const int topEnd = rd.s8(pos + 2);
					if (chunks < 0)
						throw std::runtime_error("VXL stream malformed: negative span length");
					for (int z = airFrom; z < topStart; z++) {
						if (z < 0 || z > 63)
							throw std::runtime_error("VXL stream malformed: air run outside 0..63");
						bits &= ~(1ull << z);
					}
					size_t cpos = pos + 4;
					for (int z = topStart; z <= topEnd; z++, cpos += 4)
						put(z, FileColorToMap(rd.u32(cpos)));
					if (topEnd == 62) // the water floor copies the voxel above it
						put(63, col[62]);
					const int topLen = topEnd - topStart + 1;
					if (chunks == 0) { // last span of the column
						pos += 4 * (size_t)(topLen < 0 ? 0 : topLen) + 4;
						break;
					}
					const int bottomLen = (chunks - 1) - topLen;
					const size_t next = pos + (size_t)chunks * 4;
					const int bottomEnd = rd.s8(next + 3); // next span's air start
					if (bottomLen < 0 || cpos + (size_t)bottomLen * 4 > next)
						throw std::runtime_error("VXL stream malformed: colour run crosses the next span header");
					pos = next;
					for (int z = bottomEnd - bottomLen; z < bottomEnd; z++, cpos += 4)
						put(z, FileColorToMap(rd.u32(cpos)));
					if (bottomEnd == 63)
						put(63, col[62]);
					airFrom = bottomEnd;
				}
				map->solid_[Idx(x, y)] = bits;
			}
		}
		return map;
	}

	std::unique_ptr<VoxelMap> VoxelMap::LoadFile(const char *path) {
		FILE *f = std::fopen(path, "rb");
		if (!f)
			throw std::runtime_error(std::string("cannot open ") + path);
		std::vector<uint8_t> buf;
		uint8_t tmp[65536];
		size_t n;
		while ((n = std::fread(tmp, 1, sizeof(tmp), f)) > 0)
			buf.insert(buf.end(), tmp, tmp + n);
		std::fclose(f);
		return Load(buf.data(), buf.size());
	}

	void VoxelMap::Set(int x, int y, int z, bool solid, uint32_t color, bool unsafe) {
		if (x < 0 || x >= DefaultWidth || y < 0 || y >= DefaultHeight || z < 0 || z >= DefaultDepth)
			throw std::invalid_argument("VoxelMap::Set: coordinate out of range");
		const uint64_t mask = 1ull << z;
		uint64_t &word = solid_[Idx(x, y)];
		bool changed = false;
		if (((word & mask) != 0) != solid) {
			word = solid ? (word | mask) : (word & ~mask);
			changed = true;
		}
		if (solid && color_[Idx(x, y) * 64 + z] != color) {
			color_[Idx(x, y) * 64 + z] = color;
			changed = true;
		}
		if (changed && !unsafe) {
			std::lock_guard<std::mutex> guard(listenersMutex_);
			for (auto *l : listeners_)
				l->GameMapChanged(x, y, z, this);
		}
	}

	void VoxelMap::AddListener(IVoxelMapListener *l) {
		std::lock_guard<std::mutex> guard(listenersMutex_);
		listeners_.push_back(l);
	}

	void VoxelMap::RemoveListener(IVoxelMapListener *l) {
		std::lock_guard<std::mutex> guard(listenersMutex_);
		auto it = std::find(listeners_.begin(), listeners_.end(), l);
		if (it != listeners_.end())
			listeners_.erase(it);
	}

	void VoxelMap::CastRay2Batch(osb200_ctx *ctx, const float *v0, const float *dir, int maxSteps,
	                             osb200_rayhit *out, size_t n) const {
		if (osb200_castray2(ctx, v0, dir, maxSteps, out, n) != 0)
			throw std::runtime_error(osb200_last_error());
	}

	void VoxelMap::CastRayBatch(osb200_ctx *ctx, const float *v0, const float *dir, float length, int32_t *out,
	                            size_t n) const {
		if (osb200_castray(ctx, v0, dir, length, out, n) != 0)
			throw std::runtime_error(osb200_last_error());
	}
} // namespace osb200
