// Single translation unit of libccm_orb_cuda.so (kernels are launched from orb_api.cu / match_api.cu; one TU keeps
// the build free of relocatable device code). Build: see build.py / __graft_entry__.build().
#include "orb_pyramid_fast.cu"
#include "orb_fast_chunks.cu"
#include "orb_quadtree.cu"
#include "orb_describe.cu"
#include "orb_api.cu"
#include "match_allpairs.cu"
#include "match_window.cu"
#include "match_bow.cu"
#include "match_wire.cu"
#include "match_api.cu"
#include "match_noalloc.cu"
