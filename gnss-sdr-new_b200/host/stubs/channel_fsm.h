/* Stand-in for src/core/receiver/channel_fsm.h (see README.md in this directory): only the events acquisition raises. SPDX-License-Identifier: GPL-3.0-or-later */
#ifndef B200_STUB_CHANNEL_FSM_H
#define B200_STUB_CHANNEL_FSM_H

class ChannelFsm
{
public:
    virtual ~ChannelFsm() = default;
    virtual bool Event_valid_acquisition() = 0;
    virtual bool Event_failed_acquisition_repeat() = 0;
    virtual bool Event_failed_acquisition_no_repeat() = 0;
};
#endif
