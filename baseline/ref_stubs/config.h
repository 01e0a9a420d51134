/* Stand-in for the autoconf-generated src/config.h of the reference (its configure.ac:97 computes the DSO name of
 * libsnappy into this macro). The reference arm runs uncompressed TeraSort, so the library is never dlopen'ed. */
#define HADOOP_SNAPPY_LIBRARY "libsnappy.so.1"
