// placeholder until the terrain pass lands
#include "ctx.h"
namespace osb {
	void terrain_destroy(osb200_ctx *) {}
} // namespace osb
