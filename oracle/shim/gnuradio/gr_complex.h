/* TEST INFRASTRUCTURE — stand-in for GNU Radio's <gnuradio/gr_complex.h> (GNU Radio is not installed here):
   the one typedef the reference's tracking libraries need from it. */
#ifndef ORACLE_SHIM_GR_COMPLEX_H
#define ORACLE_SHIM_GR_COMPLEX_H
#include <complex>
typedef std::complex<float> gr_complex;
typedef std::complex<double> gr_complexd;
#endif
