/* Stand-in for the CMake-generated OpenSpades.h (template: OpenSpades.h.in). Test infrastructure only. */
#pragma once
#define OpenSpades_VERSION_MAJOR 0
#define OpenSpades_VERSION_MINOR 1
#define OpenSpades_VERSION_REVISION 5
#define OPENSPADES_EXTRATITLE ""
#define OpenSpades_VER_STR "OpenSpades (oracle build)"
#define PACKAGE_STRING OpenSpades_VER_STR
#define OPENSPADES_COMPILER_STR "g++"
#define SDL_VERSION_STR "shim"
#define ZLIB_VERSION_STR "none"
#define CURL_VERSION_STR "none"
#define RESDIR "/nonexistent"
#define CMAKE_INSTALL_PREFIX "/nonexistent"
#define OPENSPADES_INSTALL_RESOURCES "/nonexistent"
