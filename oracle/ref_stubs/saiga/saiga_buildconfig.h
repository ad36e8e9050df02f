// Stand-in for the header the reference's CMake would generate; lets a few of its translation units compile
// standalone for oracle/_ref (test infrastructure).
#pragma once
#define SAIGA_USE_EIGEN
#define SAIGA_VERSION_MAJOR 1
#define SAIGA_VERSION_MINOR 3
#define SAIGA_VERSION_PATCH 2
#define SAIGA_PROJECT_SOURCE_DIR ""
#define SAIGA_INSTALL_PREFIX ""
#define SAIGA_BUILD_INCLUDE_PREFIX ""
#define SAIGA_SHADER_PREFIX ""
#define SAIGA_INCLUDE_INSTALL_PREFIX ""
