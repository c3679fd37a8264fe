// oracle/ref_neighbourhoods_main.cpp -- TEST INFRASTRUCTURE.
// Command-line front end for the reference's neibourhoods.cpp, compiled from the source where it lies
// under /root/reference (oracle/Makefile passes its directory with -I; nothing is copied).  The
// reference's own main() has its file names hard-wired (neibourhoods.cpp:101-124), so it is renamed
// away and read_do_and_write() (neibourhoods.cpp:57-99) is called with the arguments given here:
//     ref_neighbourhoods <sourcefile> <outfile> <radius> <classic 0|1>
#define main ref_main_unused
#include "neibourhoods.cpp"
#undef main
#include <cstdlib>

int main(int argc, char** argv) {
    if (argc != 5) return 2;
    return read_do_and_write(argv[1], argv[2], (size_t)atoi(argv[3]), atoi(argv[4]) != 0);
}
