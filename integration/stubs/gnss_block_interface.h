// Stand-in for src/core/interfaces/gnss_block_interface.h:57-73.
#ifndef B200_STUB_GNSS_BLOCK_INTERFACE_H
#define B200_STUB_GNSS_BLOCK_INTERFACE_H
#include <gnuradio/block.h>
#include <memory>
#include <string>
#include <utility>

template <typename T>
using gnss_shared_ptr = std::shared_ptr<T>;
template <typename C, typename... Args>
gnss_shared_ptr<C> gnss_make_shared(Args&&... args)
{
    return std::make_shared<C>(std::forward<Args>(args)...);
}

class GNSSBlockInterface
{
public:
    virtual ~GNSSBlockInterface() = default;
    virtual std::string role() = 0;
    virtual std::string implementation() = 0;
    virtual size_t item_size() = 0;
    virtual void connect(gr::top_block_sptr top_block) = 0;
    virtual void disconnect(gr::top_block_sptr top_block) = 0;
    virtual gr::basic_block_sptr get_left_block() = 0;
    virtual gr::basic_block_sptr get_right_block() = 0;
};
#endif
