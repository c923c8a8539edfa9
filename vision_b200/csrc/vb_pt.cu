#include "vb_internal.h"
void vb_pt_destroy(vb_ctx *) {}
