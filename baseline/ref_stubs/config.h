/* stand-in for the autoconf-generated src/config.h */
