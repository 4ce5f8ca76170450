/* Hand-written stand-in for the generated vcs_version.h (cmake/dav1d.cmake generates it). */
#define DAV1D_VERSION "1.5.3-oracle"
