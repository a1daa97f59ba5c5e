/* placeholder main until the headless reference GPU driver lands (see oracle/Makefile ref_gpu) */
int main() { return 0; }
