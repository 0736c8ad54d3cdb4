// oracle/compat/ccutils/easyargs.hpp — TEST INFRASTRUCTURE (oracle).
// Re-creation of the X-macro CLI generator the reference includes after defining
// REQUIRED_ARGS / OPTIONAL_ARGS / BOOLEAN_ARGS (dp.cpp:108-124, fsdp.cpp:165-182):
// args_t, make_default_args(), parse_args(), print_help(). Required arguments are
// positional in declaration order; optional ones take a value after their flag.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifndef REQUIRED_ARGS
#define REQUIRED_ARGS
#endif
#ifndef OPTIONAL_ARGS
#define OPTIONAL_ARGS
#endif
#ifndef BOOLEAN_ARGS
#define BOOLEAN_ARGS
#endif

typedef struct {
#define REQUIRED_STRING_ARG(name, label, desc) char *name;
#define REQUIRED_INT_ARG(name, label, desc) int name;
#define OPTIONAL_STRING_ARG(name, def, flag, label, desc) char *name;
#define OPTIONAL_INT_ARG(name, def, flag, label, desc) int name;
#define BOOLEAN_ARG(name, flag, desc) bool name;
    REQUIRED_ARGS OPTIONAL_ARGS BOOLEAN_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
#undef OPTIONAL_STRING_ARG
#undef OPTIONAL_INT_ARG
#undef BOOLEAN_ARG
} args_t;

static inline args_t make_default_args(void) {
    args_t a;
    memset(&a, 0, sizeof(a));
#define REQUIRED_STRING_ARG(name, label, desc) a.name = NULL;
#define REQUIRED_INT_ARG(name, label, desc) a.name = 0;
#define OPTIONAL_STRING_ARG(name, def, flag, label, desc) a.name = def;
#define OPTIONAL_INT_ARG(name, def, flag, label, desc) a.name = def;
#define BOOLEAN_ARG(name, flag, desc) a.name = false;
    REQUIRED_ARGS OPTIONAL_ARGS BOOLEAN_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
#undef OPTIONAL_STRING_ARG
#undef OPTIONAL_INT_ARG
#undef BOOLEAN_ARG
    return a;
}

static inline bool parse_args(int argc, char *argv[], args_t *a) {
    int npos = 0, want = 0;
#define REQUIRED_STRING_ARG(name, label, desc) want++;
#define REQUIRED_INT_ARG(name, label, desc) want++;
    REQUIRED_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
    for (int i = 1; i < argc; ++i) {
        bool matched = false;
#define OPTIONAL_STRING_ARG(name, def, flag, label, desc)                                          \
    if (!matched && !strcmp(argv[i], flag)) { if (i + 1 >= argc) return false; a->name = argv[++i]; matched = true; }
#define OPTIONAL_INT_ARG(name, def, flag, label, desc)                                             \
    if (!matched && !strcmp(argv[i], flag)) { if (i + 1 >= argc) return false; a->name = atoi(argv[++i]); matched = true; }
#define BOOLEAN_ARG(name, flag, desc)                                                              \
    if (!matched && !strcmp(argv[i], flag)) { a->name = true; matched = true; }
        OPTIONAL_ARGS BOOLEAN_ARGS
#undef OPTIONAL_STRING_ARG
#undef OPTIONAL_INT_ARG
#undef BOOLEAN_ARG
        if (matched) continue;
        if (argv[i][0] == '-' && argv[i][1] && !(argv[i][1] >= '0' && argv[i][1] <= '9')) return false; /* unknown flag */
        int k = 0;
#define REQUIRED_STRING_ARG(name, label, desc) if (k++ == npos) a->name = argv[i];
#define REQUIRED_INT_ARG(name, label, desc) if (k++ == npos) a->name = atoi(argv[i]);
        REQUIRED_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
        npos++;
    }
#define BOOLEAN_ARG(name, flag, desc) if (!strcmp(flag, "-h") && a->name) return true;
    BOOLEAN_ARGS
#undef BOOLEAN_ARG
    return npos == want;
}

static inline void print_help(const char *prog) {
    printf("Usage: %s", prog);
#define REQUIRED_STRING_ARG(name, label, desc) printf(" <%s>", label);
#define REQUIRED_INT_ARG(name, label, desc) printf(" <%s>", label);
    REQUIRED_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
    printf(" [options]\n\nArguments:\n");
#define REQUIRED_STRING_ARG(name, label, desc) printf("  %-24s %s\n", label, desc);
#define REQUIRED_INT_ARG(name, label, desc) printf("  %-24s %s\n", label, desc);
    REQUIRED_ARGS
#undef REQUIRED_STRING_ARG
#undef REQUIRED_INT_ARG
    printf("\nOptions:\n");
#define OPTIONAL_STRING_ARG(name, def, flag, label, desc) printf("  %-4s <%-16s> %s (default: \"%s\")\n", flag, label, desc, (const char *)def);
#define OPTIONAL_INT_ARG(name, def, flag, label, desc) printf("  %-4s <%-16s> %s (default: %d)\n", flag, label, desc, (int)def);
#define BOOLEAN_ARG(name, flag, desc) printf("  %-24s %s\n", flag, desc);
    OPTIONAL_ARGS BOOLEAN_ARGS
#undef OPTIONAL_STRING_ARG
#undef OPTIONAL_INT_ARG
#undef BOOLEAN_ARG
}
