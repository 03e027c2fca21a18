/* TEST INFRASTRUCTURE ONLY — fasta_handler.c includes htslib/hfile.h but uses nothing from it. */
