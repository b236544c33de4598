// mps.cpp -- the `mps` executable: a shell around libcusk_b200.so (see mpsmain.cu for the contract).
extern "C" int cusk_mps_main(int argc, char **argv);
int main(int argc, char **argv) { return cusk_mps_main(argc, argv); }
