/* ref_camera_shim.cpp — the reference's camera-model member bodies, compiled where they lie (cut into _ref/*.inc by oracle/Makefile).
 * TEST INFRASTRUCTURE ONLY.
 *
 * A translation unit of its own, and on purpose without `using namespace std`: src/CameraModels/KannalaBrandt8.cpp has none either
 * (nor have the headers it includes), so its unqualified `cos(psi)` / `sin(psi)` on a float (:39-40) are the C library's double
 * functions and the pixel coordinate is evaluated in double and rounded once; with std's overloads visible the same text would call
 * cosf / sinf.  The check below pins which one this build takes. */
#include <cmath>
#include <type_traits>

#include "ref_camera.h"

namespace ORB_SLAM3 {
static_assert(std::is_same<decltype(cos(1.0f)), double>::value, "cos(float) must resolve to ::cos(double) as in the reference's own translation unit");
#include "_ref/Pinhole_members.inc"
#include "_ref/KannalaBrandt8_members.inc"
}  // namespace ORB_SLAM3
