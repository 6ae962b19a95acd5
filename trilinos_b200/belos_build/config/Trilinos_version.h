#ifndef TRILINOS_VERSION_H
#define TRILINOS_VERSION_H
/* reference tree: Version.cmake:64-67 -> 12.17 (Dev) */
#define TRILINOS_MAJOR_VERSION 12
#define TRILINOS_MAJOR_MINOR_VERSION 121700
#define TRILINOS_VERSION_STRING "12.17 (Dev)"
#endif
