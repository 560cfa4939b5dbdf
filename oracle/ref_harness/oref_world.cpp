// World::ClusterizeBlocks is a file-static function of Client/World.cpp: the unmodified file is compiled as part of
// this translation unit so that the harness can call it.
#include <Client/World.cpp>
#include <cstdint>
#include <map>
#include <tuple>
extern "C" __attribute__((visibility("default")))
int oref_clusterize_blocks(const int16_t *cells, int n, int32_t *cluster) {
	std::vector<spades::client::CellPos> in;
	for (int i = 0; i < n; i++)
		in.emplace_back(cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]);
	auto res = spades::client::ClusterizeBlocks(in);
	// cluster[i] = index of the returned vector that holds cell i (a cell named twice gets the same index twice)
	std::map<std::tuple<int, int, int>, std::vector<int>> where;
	for (int i = 0; i < n; i++)
		where[{cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]}].push_back(i);
	int k = 0;
	for (auto &v : res) {
		for (auto &p : v)
			for (int i : where[{p.x, p.y, p.z}])
				cluster[i] = k;
		k++;
	}
	return k;
}
