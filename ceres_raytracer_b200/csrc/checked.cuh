// checked.cuh -- index checks of the -DCERES_CHECKED build (included first by ceres_b200.cu).
//
// Checked build (-DCERES_CHECKED, libceres_b200_checked.so; profiles/run_sanitizers.sh): every data-dependent index of the
// kernels is compared with its bound before it is used; a violation sets a CHECK_* bit in the scene's error word and
// the index is replaced by 0, so the run continues and the host reports it (ceres_scene_frame_error).  compute-sanitizer
// is not available on the GPU pool; this is the memcheck of this repository.
#define CHECK_PAIR_INDEX   0x0100ull     // node-pair index >= pairs in the scene
#define CHECK_SLOT_INDEX   0x0200ull     // triangle slot >= triangles in the scene
#define CHECK_STACK_DEPTH  0x0400ull     // traversal stack pushed beyond the scene's depth
#define CHECK_QUEUE_INDEX  0x0800ull     // record queue written / read outside the launch's region
#define CHECK_READY_INDEX  0x1000ull     // ready[] touched outside the launch's region
#define CHECK_PIXEL_INDEX  0x2000ull     // pixel / id stored outside the launch's rows
#define CHECK_STACK_UNDER  0x4000ull     // traversal stack popped below its bottom
#ifdef CERES_CHECKED
#define CERES_CLAMP(idx, limit, bit, sc) do { if (!((idx) < (limit))) { if ((sc).check) atomicOr((sc).check, (unsigned long long)(bit)); (idx) = 0; } } while (0)
#define CERES_CK(cond, bit, sc) do { if (!(cond) && (sc).check) atomicOr((sc).check, (unsigned long long)(bit)); } while (0)
#else
#define CERES_CLAMP(idx, limit, bit, sc) do { } while (0)
#define CERES_CK(cond, bit, sc) do { } while (0)
#endif

