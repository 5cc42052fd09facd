// Stand-in for <gnuradio/block.h>, <gnuradio/io_signature.h>, <gnuradio/top_block.h> (GNU Radio is not installed here).
#ifndef B200_STUB_GNURADIO_BLOCK_H
#define B200_STUB_GNURADIO_BLOCK_H
#include <pmt/pmt.h>
#include <cstddef>
#include <functional>
#include <memory>
#include <string>
#include <utility>
#include <vector>

using gr_vector_int = std::vector<int>;
using gr_vector_const_void_star = std::vector<const void*>;
using gr_vector_void_star = std::vector<void*>;

namespace gr
{
class io_signature
{
public:
    using sptr = std::shared_ptr<io_signature>;
    static sptr make(int min_streams, int max_streams, int sizeof_stream_item) { return sptr(new io_signature{min_streams, max_streams, sizeof_stream_item}); }
    int min_streams, max_streams, sizeof_stream_item;
};

class basic_block
{
public:
    explicit basic_block(std::string name) : d_name(std::move(name)) {}
    virtual ~basic_block() = default;
    const std::string& name() const { return d_name; }
    long unique_id() const { return 0; }
    void message_port_register_out(const pmt::pmt_t& port) { (void)port; }
    // the test harness listens here instead of a connected message port
    std::function<void(const pmt::pmt_t&, const pmt::pmt_t&)> on_message_port_pub;
    void message_port_pub(const pmt::pmt_t& port, const pmt::pmt_t& msg)
    {
        if (on_message_port_pub) on_message_port_pub(port, msg);
    }

private:
    std::string d_name;
};
using basic_block_sptr = std::shared_ptr<basic_block>;

class block : public basic_block
{
public:
    block(const std::string& name, io_signature::sptr in, io_signature::sptr out) : basic_block(name), d_in(std::move(in)), d_out(std::move(out)) {}
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    virtual void forecast(int noutput_items, gr_vector_int& ninput_items_required)
    {
        for (auto& n : ninput_items_required) n = noutput_items;
    }
    virtual bool start() { return true; }
    void consume_each(int how_many_items) { d_consumed += how_many_items; }
    long consumed_so_far() const { return d_consumed; }  // harness only
    io_signature::sptr input_signature() const { return d_in; }
    io_signature::sptr output_signature() const { return d_out; }

private:
    io_signature::sptr d_in, d_out;
    long d_consumed{0};
};

class top_block
{
public:
    void connect(const basic_block_sptr&, int, const basic_block_sptr&, int) {}
    void disconnect(const basic_block_sptr&, int, const basic_block_sptr&, int) {}
};
using top_block_sptr = std::shared_ptr<top_block>;
}  // namespace gr
#endif
