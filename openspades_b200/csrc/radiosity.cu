// placeholder until the radiosity pass lands
#include "ctx.h"
namespace osb {
	void radiosity_destroy(osb200_ctx *) {}
	void radiosity_map_replaced(osb200_ctx *) {}
	void radiosity_mark_dirty(osb200_ctx *, int, int, int) {}
} // namespace osb
