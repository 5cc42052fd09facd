/*
 * pcps_acquisition_b200_block.cc — see the header.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "pcps_acquisition_b200_block.h"
#include "gnss_synchro.h"
#include <pmt/pmt.h>

pcps_acquisition_b200_block_sptr pcps_make_acquisition_b200_block(const Acq_Conf& conf_)
{
    return pcps_acquisition_b200_block_sptr(new pcps_acquisition_b200_block(conf_));
}


// Same stream signature as the reference block (pcps_acquisition.cc:50-52): one input of the configured item type, an
// optional Gnss_Synchro output for the acquisition monitor; decisions leave through the "events" message port (:83).
pcps_acquisition_b200_block::pcps_acquisition_b200_block(const Acq_Conf& conf_)
    : gr::block("pcps_acquisition_b200",
          gr::io_signature::make(1, 1, static_cast<int>(conf_.it_size)),
          gr::io_signature::make(0, 1, sizeof(Gnss_Synchro))),
      d_core(pcps_make_acquisition_b200(conf_))
{
    this->message_port_register_out(pmt::mp("events"));
    d_core->consume_each = [this](int n_items) { this->consume_each(n_items); };
    d_core->publish_event = [this](long event) { this->message_port_pub(pmt::mp("events"), pmt::from_long(event)); };
}


// The block is a sink for the scheduler's purposes: any number of available items is worth a call (the reference block does
// not override forecast either way; one item is what gr::block's default asks for with a fixed rate of 1).
void pcps_acquisition_b200_block::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    (void)noutput_items;
    for (auto& n : ninput_items_required) n = 1;
}


int pcps_acquisition_b200_block::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
    gr_vector_void_star& output_items)
{
    return d_core->general_work(noutput_items, ninput_items, input_items, output_items);
}
