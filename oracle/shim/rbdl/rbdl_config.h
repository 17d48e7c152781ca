// TEST INFRASTRUCTURE (oracle build only) -- hand-written stand-in for the cmake-generated
// rbdl_config.h (template: external/rbdl-v2.5.0/include/rbdl/rbdl_config.h.cmake:1-74).
// Selects RBDL's bundled SimpleMath backend because Eigen3 is not installed in this image.
#ifndef RBDL_CONFIG_H
#define RBDL_CONFIG_H
#define RBDL_API_VERSION ((2 << 16) + (5 << 8) + 0)
#define RBDL_USE_SIMPLE_MATH
#define RBDL_BUILD_REVISION "unknown"
#define RBDL_BUILD_TYPE "unknown"
#define RBDL_BUILD_BRANCH "unknown"
#define RBDL_BUILD_STATIC
#define RBDL_DLLAPI
#define RBDL_LOCAL
#endif
