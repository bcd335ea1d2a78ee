/* placeholder so that `make oracle` has a translation unit; replaced below */
int cm_oracle_version(void) { return 1; }
