#include <ompl/base/State.h>
int main() { return 0; }
