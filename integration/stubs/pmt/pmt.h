// Stand-in for <pmt/pmt.h>: the two calls the acquisition block makes (port name, event number).
#ifndef B200_STUB_PMT_H
#define B200_STUB_PMT_H
#include <memory>
#include <string>
namespace pmt
{
struct pmt_base
{
    std::string symbol;
    long value{0};
};
using pmt_t = std::shared_ptr<pmt_base>;
inline pmt_t mp(const std::string& s)
{
    auto p = std::make_shared<pmt_base>();
    p->symbol = s;
    return p;
}
inline pmt_t from_long(long v)
{
    auto p = std::make_shared<pmt_base>();
    p->value = v;
    return p;
}
inline long to_long(const pmt_t& p) { return p->value; }
}  // namespace pmt
#endif
