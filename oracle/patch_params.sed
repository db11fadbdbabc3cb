# parameters.h:16 -- the one-token fix g++ needs (SURVEY.md D8)
s/^#define STCONST static const[[:space:]]*$/#define STCONST static constexpr/
# parameters.h:77 -- let a config switch the intracellular-throw flag off
s/^([[:space:]]*)#define AVOID_INTRACELLULAR_THROW[[:space:]]*$/\1#ifndef REF_NO_AVOID_INTRACELLULAR_THROW\n\1#define AVOID_INTRACELLULAR_THROW\n\1#endif/
# every knob a BASELINE.json config varies: keep the shipped initialiser as the default
s/^([[:space:]]*STCONST[[:space:]]+(int|double)[[:space:]]+)(num_water|num_cells|cell_r|mmoment|bound|water_start_bound|hashDim|maxNeighbors|border|scale|sprintSteps|totaltime|num_threads|taucp)([[:space:]]*=[[:space:]]*)([^;]+);(.*)$/#ifdef REF_\3\n\1\3\4REF_\3;\6\n#else\n\1\3\4\5;\6\n#endif/
# parameters.h:116-117 -- BASELINE config 3: activate the reference's own commented low-permeability formulas
s/^([[:space:]]*)reflectIO = 0; \/\/ (1 - sqrt.*;)[[:space:]]*$/#ifdef REF_LOW_PERMEABILITY\n\1reflectIO = \2\n#else\n\1reflectIO = 0;\n#endif/
s/^([[:space:]]*)reflectOI = 0; \/\/ (1 - \(\(1 - reflectIO\).*;)[[:space:]]*$/#ifdef REF_LOW_PERMEABILITY\n\1reflectOI = \2\n#else\n\1reflectOI = 0;\n#endif/
# parameters.h:55-56 -- let a config select CONSTANT_KICK instead of the shipped RANDOM_KICK
s/^([[:space:]]*)#undef CONSTANT_KICK[[:space:]]*$/\1#ifdef REF_CONSTANT_KICK\n\1#define CONSTANT_KICK\n\1#endif/
s/^([[:space:]]*)#define RANDOM_KICK[[:space:]]*$/\1#ifndef REF_CONSTANT_KICK\n\1#define RANDOM_KICK\n\1#endif/
