// Stand-in for the two conversion blocks the reference adapters put in front of the acquisition block for "cbyte" items
// (src/algorithms/libs/complex_byte_to_float_x2.h, <gnuradio/blocks/float_to_complex.h>).
#ifndef B200_STUB_CONVERSION_BLOCKS_H
#define B200_STUB_CONVERSION_BLOCKS_H
#include <gnuradio/block.h>
#include <memory>
using complex_byte_to_float_x2_sptr = gr::basic_block_sptr;
inline complex_byte_to_float_x2_sptr make_complex_byte_to_float_x2() { return std::make_shared<gr::basic_block>("complex_byte_to_float_x2"); }
namespace gr
{
namespace blocks
{
struct float_to_complex
{
    using sptr = gr::basic_block_sptr;
    static sptr make() { return std::make_shared<gr::basic_block>("float_to_complex"); }
};
}  // namespace blocks
}  // namespace gr
#endif
