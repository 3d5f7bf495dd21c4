/* Test-infrastructure shim (oracle/): boost is not installed in this image and the
 * reference's mask.h:5-8 only needs BOOST_PP_DIV (mask.h:15-25). Not product code. */
#ifndef BOOST_PP_DIV
#define BOOST_PP_DIV(a, b) ((a) / (b))
#endif
