// oracle shim: stands in for the CMake-generated config header
// (reference CMake/amrVolumeRendererConfig.h.in).
#ifndef AMRVOLUMERENDERERCONFIG_H
#define AMRVOLUMERENDERERCONFIG_H
#define AMRVOLUMERENDERER_APP_NAME "oracle"
#endif
