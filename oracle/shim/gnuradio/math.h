/* TEST INFRASTRUCTURE — stand-in for GNU Radio's <gnuradio/math.h>. gr::fast_atan2f is a table-driven atan2 of the
   absent third-party library; it is only reached from pll_four_quadrant_atan / fll_four_quadrant_atan, which the
   compared path (Costas loop, fll_diff_atan) never calls. Mapped to atan2f so that the file links. */
#ifndef ORACLE_SHIM_GR_MATH_H
#define ORACLE_SHIM_GR_MATH_H
#include <cmath>
namespace gr
{
inline float fast_atan2f(float y, float x) { return std::atan2(y, x); }
}  // namespace gr
#endif
