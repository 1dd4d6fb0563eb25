// tiled.cu -- tiled fast-path step kernel (placeholder until the tiled layout lands: reports "unsupported",
// so every configuration runs through the generic CSR-gather kernel in csr_step.cu).
#include "engine.h"

namespace tb {

bool tiled_supported(const Engine *) { return false; }
int tiled_prepare(Engine *) { return TB_OK; }
int launch_tiled_step(Engine *e, bool, bool) { return fail(e, TB_ERR_UNSUPPORTED, "tiled kernel not built"); }
int tiled_to_canonical(Engine *e) { e->canon_stale = false; return TB_OK; }
void tiled_destroy(Engine *) {}

}  // namespace tb
