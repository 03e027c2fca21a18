/* TEST INFRASTRUCTURE ONLY — see htslib/faidx.h in this directory.  Reads the whole FASTA into memory. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>
#include "htslib/faidx.h"

struct sa_faidx_stub {
    int n;
    char **name;
    char **seq;
    long *len;
};

faidx_t *fai_load(const char *fn) {
    FILE *f = fopen(fn, "r");
    if (!f) return NULL;
    faidx_t *fa = calloc(1, sizeof(*fa));
    char *line = NULL;
    size_t cap = 0;
    ssize_t got;
    long scap = 0;
    while ((got = getline(&line, &cap, f)) >= 0) {
        while (got > 0 && (line[got - 1] == '\n' || line[got - 1] == '\r')) line[--got] = 0;
        if (line[0] == '>') {
            fa->n++;
            fa->name = realloc(fa->name, sizeof(char *) * fa->n);
            fa->seq = realloc(fa->seq, sizeof(char *) * fa->n);
            fa->len = realloc(fa->len, sizeof(long) * fa->n);
            char *e = line + 1;
            while (*e && !isspace((unsigned char) *e)) e++;
            *e = 0;
            fa->name[fa->n - 1] = strdup(line + 1);
            fa->seq[fa->n - 1] = NULL;
            fa->len[fa->n - 1] = 0;
            scap = 0;
        } else if (fa->n > 0 && got > 0) {
            int i = fa->n - 1;
            if (fa->len[i] + got + 1 > scap) {
                scap = 2 * (fa->len[i] + got + 1);
                fa->seq[i] = realloc(fa->seq[i], scap);
            }
            memcpy(fa->seq[i] + fa->len[i], line, got);
            fa->len[i] += got;
            fa->seq[i][fa->len[i]] = 0;
        }
    }
    free(line);
    fclose(f);
    return fa;
}

char *faidx_fetch_seq(const faidx_t *fa, const char *c_name, int beg, int end, int *len) {
    for (int i = 0; fa && i < fa->n; i++) {
        if (strcmp(fa->name[i], c_name) != 0) continue;
        if (beg < 0) beg = 0;
        if (end >= fa->len[i]) end = (int) fa->len[i] - 1;
        int n = end >= beg ? end - beg + 1 : 0;
        char *s = malloc(n + 1);
        if (n) memcpy(s, fa->seq[i] + beg, n);
        s[n] = 0;
        *len = n;
        return s;
    }
    *len = -2;
    return NULL;
}

void fai_destroy(faidx_t *fa) {
    if (!fa) return;
    for (int i = 0; i < fa->n; i++) { free(fa->name[i]); free(fa->seq[i]); }
    free(fa->name); free(fa->seq); free(fa->len); free(fa);
}
