/* Hand-written stand-in for the cmake-generated Triutils_config.h (oracle/_ref only: the reference's
 * Harwell-Boeing reader is compiled as the parity oracle of b200_read_hb). */
#ifndef TRIUTILS_CONFIG_H
#define TRIUTILS_CONFIG_H
#define HAVE_CMATH
#define HAVE_CSTDIO
#define HAVE_CSTDLIB
#define HAVE_IOSTREAM
#define HAVE_STRING
#define HAVE_VECTOR
#define HAVE_MAP
#define HAVE_STDEXCEPT
#endif
