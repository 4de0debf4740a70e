// Same three-member surface as the reference's Sequencer_module (src/sequence.h:5,
// src/pimpl.h:18-28), so src/tksm.cpp:164-166 compiles unchanged against this build.
#ifndef _SEQUENCER_H
#define _SEQUENCER_H
#include <memory>

class Sequencer_module {
    class impl;
    std::unique_ptr<impl> pimpl;

public:
    Sequencer_module(int argc, char **argv);
    ~Sequencer_module();
    int run();
};

#endif
