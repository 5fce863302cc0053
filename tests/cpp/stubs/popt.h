/* Declarations-only stand-in for <popt.h> (absent in this image): just enough of the public popt interface for
 * `g++ -fsyntax-only` of the reference's sdrx.cpp against the drop-in headers (tests/test_dropin_syntax.py). */
#ifndef STUB_POPT_H
#define STUB_POPT_H
#ifdef __cplusplus
extern "C" {
#endif
#define POPT_ARG_NONE 0U
#define POPT_ARG_STRING 1U
#define POPT_ARG_INT 2U
#define POPT_ARG_FLOAT 8U
#define POPT_CONTEXT_POSIXMEHARDER (1U << 2)
#define POPT_ERROR_NOARG -10
#define POPT_ERROR_BADOPT -11
#define POPT_ERROR_BADNUMBER -17
#define POPT_ERROR_OVERFLOW -18
#define POPT_TABLEEND { NULL, '\0', 0, NULL, 0, NULL, NULL }
struct poptOption {
    const char *longName; char shortName; unsigned int argInfo; void *arg; int val;
    const char *descrip; const char *argDescrip;
};
typedef struct poptContext_s *poptContext;
poptContext poptGetContext(const char *name, int argc, const char **argv, const struct poptOption *options, unsigned int flags);
poptContext poptFreeContext(poptContext con);
int poptGetNextOpt(poptContext con);
const char *poptGetArg(poptContext con);
const char *poptPeekArg(poptContext con);
const char *poptBadOption(poptContext con, unsigned int flags);
const char *poptStrerror(const int error);
void poptPrintHelp(poptContext con, FILE *fp, int flags);
void poptPrintUsage(poptContext con, FILE *fp, int flags);
void poptSetOtherOptionHelp(poptContext con, const char *text);
#ifdef __cplusplus
}
#endif
#endif
