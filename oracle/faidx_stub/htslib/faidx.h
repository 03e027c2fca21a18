/*
 * TEST INFRASTRUCTURE ONLY — a plain-FASTA stand-in for the three htslib faidx calls the reference's
 * impl/fasta_handler.c makes (fai_load, faidx_fetch_seq, fai_destroy; /root/reference/impl/fasta_handler.c:16-44).
 * htslib is not installed in this image, so the UNMODIFIED fasta_handler.c and signalMachine.c are compiled against this
 * header (oracle/Makefile, target cli) to link the reference's real CLI twice: CPU objects / GPU shim objects.
 */
#ifndef SA_FAIDX_STUB_H
#define SA_FAIDX_STUB_H
typedef struct sa_faidx_stub faidx_t;
faidx_t *fai_load(const char *fn);
/* [p_beg_i, p_end_i] 0-based inclusive, clipped to the record; *len = -2 when the record does not exist */
char *faidx_fetch_seq(const faidx_t *fai, const char *c_name, int p_beg_i, int p_end_i, int *len);
void fai_destroy(faidx_t *fai);
#endif
