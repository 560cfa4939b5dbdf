/*
 * TEST INFRASTRUCTURE ONLY.  Three symbols the reference's Core/ objects need at link time but
 * whose real definitions drag in FLTK-preference import and jsoncpp.
 */
#include <Core/Debug.h>
#include <Core/FltkPreferenceImporter.h>
#include <Core/VoxelModelLoader.h>

namespace spades {
	std::vector<std::pair<std::string, std::string>> ImportFltkPreference() { return {}; }
	void DeleteFltkPreference() {}
	Handle<VoxelModel> VoxelModelLoader::Load(const char *) {
		SPRaise("VoxelModelLoader is not available in the oracle build");
	}
} // namespace spades
