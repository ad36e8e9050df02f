// Stand-in for the reference's generated module list (see saiga_buildconfig.h here).
#pragma once
#define SAIGA_CORE
#define SAIGA_VISION
