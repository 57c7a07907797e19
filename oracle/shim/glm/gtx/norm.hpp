// TEST INFRASTRUCTURE — stand-in for <glm/gtx/norm.hpp> (see glm/glm.hpp in this directory).
// length2(v) = dot(v, v), as in GLM.  Used by the reference at include/light_impl.h:36.
#ifndef RT_ORACLE_SHIM_GLM_NORM_HPP
#define RT_ORACLE_SHIM_GLM_NORM_HPP
#include "../glm.hpp"
namespace glm {
template <typename T, qualifier Q>
T length2(const vec<3, T, Q> &v) { return dot(v, v); }
}
#endif
