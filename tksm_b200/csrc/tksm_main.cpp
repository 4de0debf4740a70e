// `tksm` front-end restricted to the Seq exit module: `tksm sequence ...` dispatches exactly like
// src/tksm.cpp:164-166 (Sequencer_module{argc - 1, argv + 1}.run()).
#include <cstdio>
#include <cstring>

#include "sequence.h"

int main(int argc, char** argv) {
    if (argc >= 2 && (!strcmp(argv[1], "sequence"))) return Sequencer_module{argc - 1, argv + 1}.run();
    if (argc >= 2 && !strcmp(argv[1], "list")) {
        printf("sequence\n");
        return 0;
    }
    fprintf(stderr, "tksm (B200 Seq build): only the `sequence` module is provided.\nusage: tksm sequence [options]\n");
    return argc >= 2 ? 1 : 0;
}
