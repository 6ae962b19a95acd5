/* Hand-written stand-in for the cmake-generated Ifpack_config.h (oracle/_ref only: the reference's
 * Ifpack_Chebyshev is compiled as the parity oracle of the Chebyshev right preconditioner).
 * No optional TPL/package is enabled: no AztecOO, no EpetraExt (point, not block, mode). */
#ifndef IFPACK_CONFIG_H
#define IFPACK_CONFIG_H
#endif
