/* unit_driver.c - TEST INFRASTRUCTURE. Runs the reference's own CuTest cases (tests/paf_unit_test.c and test_paf of
 * tests/paf_test.c, both compiled UNCHANGED from /root/reference) against libpaffy_b200.so + libpaffy_b200_stpaf.a, one
 * case at a time, printing "PASS name" / "FAIL name: message". The shell-script cases of paf_test.c need lastz and the
 * network and are never run. Arguments: substrings of case names to skip (the three paf_encode_mismatches cases need a
 * CUDA device: the CPU test skips them, the GPU test runs everything). Exit code = number of failures.
 * Must be started from a directory that holds ./tests/human_chimp.paf (tests/paf_test.c:9 uses a relative path). */
#include <stdio.h>
#include <string.h>

#include "CuTest.h"

CuSuite *addPafTestSuite(void);
CuSuite *addPafUnitTestSuite(void);

static int run_suite(CuSuite *s, const char *only, int n_skip, char **skip) {
    int failures = 0;
    for (int i = 0; i < s->count; i++) {
        CuTest *t = s->list[i];
        int skipped = only && strcmp(t->name, only) != 0;
        for (int k = 0; k < n_skip && !skipped; k++) skipped = strstr(t->name, skip[k]) != NULL;
        if (skipped) {
            if (!only) printf("SKIP %s\n", t->name);
            continue;
        }
        CuTestRun(t);
        if (t->failed) {
            failures++;
            printf("FAIL %s: %s\n", t->name, t->message ? t->message : "");
        } else {
            printf("PASS %s\n", t->name);
        }
        fflush(stdout);
    }
    return failures;
}

int main(int argc, char **argv) {
    int failures = run_suite(addPafUnitTestSuite(), NULL, argc - 1, argv + 1);
    failures += run_suite(addPafTestSuite(), "test_paf", 0, NULL);
    printf("failures: %d\n", failures);
    return failures;
}
