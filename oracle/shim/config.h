/* Hand-written stand-in for the autotools-generated config.h (autotools are absent on this image).
 * Test infrastructure only: used when compiling the UNMODIFIED reference sources from /root/reference
 * into oracle/_ref/.  Nothing under oracle/ is on the product path. */
#ifndef TJB_ORACLE_SHIM_CONFIG_H
#define TJB_ORACLE_SHIM_CONFIG_H
#define HAVE_ZLIB 1
#define PACKAGE "tatajuba"
#define PACKAGE_STRING "tatajuba 1.0.5"
#define PACKAGE_VERSION "1.0.5"
#ifndef _GNU_SOURCE
#define _GNU_SOURCE 1
#endif
#define SIZEOF_INT 4
#define SIZEOF_DOUBLE 8
#define SIZEOF_LONG_LONG_INT 8
#define HAVE_UNISTD_H 1
#define HAVE_SYS_TIME_H 1
#define HAVE_SYS_TIMES_H 1
#endif
