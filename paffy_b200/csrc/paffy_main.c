/* paffy_b200 - command line front end for the subcommands libpaffy_b200.so implements, dispatched by name the way the
 * reference's paffy_main.c:52-77 does ("paffy_b200 add_mismatches ..." == "paffy add_mismatches ..."). */
#include <stdio.h>
#include <string.h>

int paffy_add_mismatches_main(int argc, char *argv[]);
int paffy_view_main(int argc, char *argv[]);
int paffy_trim_main(int argc, char *argv[]);

int main(int argc, char *argv[]) {
    if (argc >= 2 && strcmp(argv[1], "add_mismatches") == 0) return paffy_add_mismatches_main(argc - 1, argv + 1);
    if (argc >= 2 && strcmp(argv[1], "view") == 0) return paffy_view_main(argc - 1, argv + 1);
    if (argc >= 2 && strcmp(argv[1], "trim") == 0) return paffy_trim_main(argc - 1, argv + 1);
    fprintf(stderr, "usage: paffy_b200 add_mismatches [fasta_file]xN [-i in.paf] [-o out.paf] [-a]\n"
                    "       paffy_b200 view [fasta_file]xN [-i in.paf] [-o out.txt] [-s] [-t] [-u X] [-v X]\n"
                    "       paffy_b200 trim [-i in.paf] [-o out.paf] [-r X] [-t X] [-f]\n");
    return argc < 2 ? 0 : 1;
}
