/* Hand-written stand-in for the generated Epetra_DLLExportMacro.h (Trilinos_DLLExportMacro.h.in
 * at the reference root): no Windows DLL decoration on Linux. */
#ifndef EPETRA_DLLEXPORTMACRO_H
#define EPETRA_DLLEXPORTMACRO_H
#define EPETRA_LIB_DLL_EXPORT
#endif
