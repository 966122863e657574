/* mg_config.h -- the lowered ("flat") game config consumed by the step kernels.
 *
 * A lowered config is one contiguous array of int32 words. It is what
 * `convert_to_cpp_game_config` (reference: packages/mettagrid/python/src/mettagrid/config/
 * mettagrid_c_config.py:30-577) produces as a tree of pybind objects, flattened so that a kernel
 * can index it. All names become dense ids at lowering time; every list whose order the
 * reference takes from a std::unordered_map is stored already in that iteration order.
 *
 * Word offsets below are in int32 units. Floats are stored as their IEEE-754 bit pattern.
 * The Python side (metta_b200/lowering.py) reads these #defines by parsing this file, so this
 * header is the single definition of the layout.
 */
#ifndef MG_CONFIG_H_
#define MG_CONFIG_H_

#define MG_CONFIG_MAGIC 0x4D473230 /* "MG20" */
#define MG_CONFIG_VERSION 4

/* ---- capacity limits (checked at lowering time, errors are loud) ---- */
#define MG_MAX_RESOURCES 13   /* <= 13 keeps libstdc++'s 13-bucket inventory map collision free */
#define MG_RES_STRIDE 16      /* storage stride of per-resource arrays */
#define MG_MAX_INV_TOKENS 8   /* digits per inventory amount (token_value_base >= 4) */
#define MG_MAX_TAGS 8         /* tags per object type */
#define MG_MAX_GROUPS 8       /* distinct agent configs (teams) */
#define MG_MAX_CELLTYPES 64   /* distinct map cell names */
#define MG_MAX_ASM_TYPES 32
#define MG_MAX_CHEST_TYPES 16
#define MG_MAX_PROTOCOLS 256
#define MG_MAX_PROTO_VIBES 8
#define MG_MAX_ACTIONS 512
#define MG_MAX_VIBES 256
#define MG_MAX_LIMITS 16      /* shared limit groups per inventory */
#define MG_MAX_REWARDS 16     /* stat reward terms per agent config */
#define MG_MAX_REGEN 8        /* regen lists per agent config */
#define MG_MAX_TRANSFERS 32   /* vibe transfer effects */
#define MG_MAX_CHEST_VIBES 16 /* vibe->delta lists per chest type */
#define MG_MAX_GOALS 16

/* ---- "item list": count, then MG_RES_STRIDE (item, value) pairs, in reference iteration order ---- */
#define MG_LIST_WORDS (1 + 2 * MG_RES_STRIDE)
#define MG_LIST_N 0
#define MG_LIST_ITEM(i) (1 + 2 * (i))
#define MG_LIST_VAL(i) (2 + 2 * (i))

/* ---- header: fixed word offsets ---- */
#define MGC_MAGIC 0
#define MGC_VERSION 1
#define MGC_TOTAL_WORDS 2
#define MGC_NUM_AGENTS 3
#define MGC_MAP_H 4
#define MGC_MAP_W 5
#define MGC_OBS_H 6
#define MGC_OBS_W 7
#define MGC_NUM_TOKENS 8
#define MGC_TOKEN_BASE 9
#define MGC_NUM_INV_TOKENS 10
#define MGC_MAX_STEPS 11
#define MGC_EPISODE_TRUNCATES 12
#define MGC_REGEN_INTERVAL 13
#define MGC_NUM_RESOURCES 14
#define MGC_NUM_ACTIONS 15
#define MGC_NUM_GROUPS 16
#define MGC_NUM_CELLTYPES 17
#define MGC_NUM_ASM_TYPES 18
#define MGC_NUM_CHEST_TYPES 19
#define MGC_NUM_PROTOCOLS 20
#define MGC_PROTOCOL_DETAILS_OBS 21
#define MGC_GOBS_ECP 22
#define MGC_GOBS_LAST_ACTION 23
#define MGC_GOBS_LAST_REWARD 24
#define MGC_GOBS_COMPASS 25
#define MGC_GOBS_GOAL 26
#define MGC_FEAT_GROUP 27
#define MGC_FEAT_FROZEN 28
#define MGC_FEAT_ECP 29
#define MGC_FEAT_LAST_ACTION 30
#define MGC_FEAT_LAST_REWARD 31
#define MGC_FEAT_VIBE 32
#define MGC_FEAT_COMPASS 33
#define MGC_FEAT_TAG 34
#define MGC_FEAT_COOLDOWN 35
#define MGC_FEAT_CLIPPED 36
#define MGC_FEAT_REMAINING_USES 37
#define MGC_FEAT_GOAL 38
#define MGC_NUM_AGENT_STATS 39   /* float slots per agent */
#define MGC_NUM_GAME_STATS 40    /* float slots per env */
#define MGC_NUM_TRANSFERS 41
#define MGC_MAX_ASSEMBLERS 42    /* per-env instance capacity (max over the map pool) */
#define MGC_MAX_CHESTS 43
#define MGC_HAS_VIBE_ASM 44      /* any assembler type whose protocol depends on neighbours */
#define MGC_NUM_INVCFG 45
/* section offsets (absolute word offsets into the blob) */
#define MGC_OFF_ACTIONS 48
#define MGC_OFF_ACTION_RES 49
#define MGC_OFF_VIBES 50
#define MGC_OFF_ATTACK 51
#define MGC_OFF_TRANSFER 52
#define MGC_OFF_GROUPS 53
#define MGC_OFF_CELLTYPES 54
#define MGC_OFF_ASM_TYPES 55
#define MGC_OFF_PROTOCOLS 56
#define MGC_OFF_CHEST_TYPES 57
#define MGC_OFF_CLIPPER 58
#define MGC_OFF_INV_FEATURES 59
#define MGC_OFF_PROTO_FEATURES 60
#define MGC_OFF_INVCFG 61
#define MGC_HEADER_WORDS 64

/* ---- ACTIONS: NUM_ACTIONS records ---- */
#define MG_ACT_WORDS 2
#define MG_ACT_KIND 0 /* MG_ACTION_* */
#define MG_ACT_ARG 1  /* move: Orientation 0..7 (N,S,W,E,NW,NE,SW,SE); change_vibe: vibe id */
#define MG_ACTION_NOOP 0
#define MG_ACTION_MOVE 1
#define MG_ACTION_CHANGE_VIBE 2

/* ---- ACTION_RES: for each of the 3 action kinds: required[16], consumed[16] ---- */
#define MG_ACTRES_WORDS (2 * MG_RES_STRIDE)
#define MG_ACTRES_REQUIRED 0
#define MG_ACTRES_CONSUMED MG_RES_STRIDE

/* ---- VIBES: MG_MAX_VIBES records: what a mover's vibe triggers (move.hpp:45-66) ---- */
#define MG_VIBE_WORDS 4
#define MG_VIBE_HANDLER 0   /* 0 none, 1 attack, 2 transfer */
#define MG_VIBE_TRANSFER 1  /* transfer effect index or -1 (transfer.hpp:88-91) */
#define MG_VIBE_BONUS 2     /* attack.vibe_bonus[vibe] (attack.hpp:166-169) */
#define MG_VIBE_RESOURCE 3  /* resource whose name equals this vibe's name, or -1 (attack.hpp:171-178) */
#define MG_HANDLER_NONE 0
#define MG_HANDLER_ATTACK 1
#define MG_HANDLER_TRANSFER 2

/* ---- ATTACK (attack.hpp:36-66) ---- */
#define MG_ATK_CONSUMED 0                              /* [16] */
#define MG_ATK_ARMOR (MG_RES_STRIDE)                   /* [16] weights */
#define MG_ATK_WEAPON (2 * MG_RES_STRIDE)              /* [16] weights */
#define MG_ATK_FREEZE (3 * MG_RES_STRIDE)
#define MG_ATK_HAS_DEFENSE (3 * MG_RES_STRIDE + 1)
#define MG_ATK_DEFENSE (3 * MG_RES_STRIDE + 2)         /* list (item, amount) */
#define MG_ATK_ACTOR_DELTA (MG_ATK_DEFENSE + MG_LIST_WORDS)
#define MG_ATK_TARGET_DELTA (MG_ATK_ACTOR_DELTA + MG_LIST_WORDS)
#define MG_ATK_LOOT (MG_ATK_TARGET_DELTA + MG_LIST_WORDS) /* list, value unused */
#define MG_ATK_WORDS (MG_ATK_LOOT + MG_LIST_WORDS)

/* ---- TRANSFER (transfer.hpp:29-39): required[16], then NUM_TRANSFERS effects ---- */
#define MG_XFER_REQUIRED 0
#define MG_XFER_EFFECTS MG_RES_STRIDE
#define MG_XFER_EFFECT_WORDS (2 * MG_LIST_WORDS)
#define MG_XFER_ACTOR 0
#define MG_XFER_TARGET MG_LIST_WORDS

/* ---- INVCFG: inventory limit config (inventory_config.hpp:14-31), shared by groups/chests ---- */
#define MG_INVCFG_ITEM_LIMIT 0                               /* [16] limit index of item or -1 */
#define MG_INVCFG_IS_MODIFIER MG_RES_STRIDE                  /* [16] */
#define MG_INVCFG_NUM_LIMITS (2 * MG_RES_STRIDE)
#define MG_INVCFG_LIMIT_BASE (2 * MG_RES_STRIDE + 1)         /* [MG_MAX_LIMITS] */
#define MG_INVCFG_LIMIT_MOD (MG_INVCFG_LIMIT_BASE + MG_MAX_LIMITS) /* [MG_MAX_LIMITS][16] bonus per item */
#define MG_INVCFG_NUM_ORDER (MG_INVCFG_LIMIT_MOD + MG_MAX_LIMITS * MG_RES_STRIDE)
#define MG_INVCFG_ORDER (MG_INVCFG_NUM_ORDER + 1)             /* [16] limited items, _limits map order (inventory.cpp:133-156) */
#define MG_INVCFG_WORDS (MG_INVCFG_ORDER + MG_RES_STRIDE)

/* ---- GROUPS: one record per distinct agent config ---- */
#define MG_GRP_GROUP_ID 0
#define MG_GRP_FREEZE_DURATION 1
#define MG_GRP_INITIAL_VIBE 2
#define MG_GRP_NUM_TAGS 3
#define MG_GRP_TAGS 4                                  /* [MG_MAX_TAGS] */
#define MG_GRP_INVCFG (MG_GRP_TAGS + MG_MAX_TAGS)      /* index into INVCFG section */
#define MG_GRP_INITIAL_INV (MG_GRP_INVCFG + 1)         /* list, in initial_inventory map order */
#define MG_GRP_NUM_REWARDS (MG_GRP_INITIAL_INV + MG_LIST_WORDS)
#define MG_GRP_REWARDS (MG_GRP_NUM_REWARDS + 1)        /* [MG_MAX_REWARDS] x MG_RWD_WORDS, stat_rewards map order */
#define MG_RWD_WORDS 6
#define MG_RWD_AGENT_STAT 0  /* agent stat slot or -1 */
#define MG_RWD_GAME_STAT 1   /* game stat slot or -1 */
#define MG_RWD_WEIGHT 2      /* f32 bits */
#define MG_RWD_HAS_MAX 3
#define MG_RWD_MAX 4         /* f32 bits */
#define MG_RWD_COLLECTIVE 5  /* f32 bits: the agent's collective's value of this stat (agent.cpp:115-117). Collective stats
                                only change through activation handlers (not lowered), so it is a constant of the config:
                                aligned.<type> member counts and collective.<res>.deposited of the initial inventory
                                (objects/collective.hpp:36-62,100-108) */
#define MG_GRP_DIVERSITY (MG_GRP_REWARDS + MG_MAX_REWARDS * MG_RWD_WORDS) /* [16] mask */
#define MG_GRP_HAS_REGEN (MG_GRP_DIVERSITY + MG_RES_STRIDE)
#define MG_GRP_NUM_REGEN (MG_GRP_HAS_REGEN + 1)
#define MG_GRP_REGEN_VIBE (MG_GRP_NUM_REGEN + 1)       /* [MG_MAX_REGEN] vibe id of each list */
#define MG_GRP_REGEN_LISTS (MG_GRP_REGEN_VIBE + MG_MAX_REGEN) /* [MG_MAX_REGEN] lists (item, delta) */
#define MG_GRP_DAMAGE_ENABLED (MG_GRP_REGEN_LISTS + MG_MAX_REGEN * MG_LIST_WORDS)
#define MG_GRP_DAMAGE_THRESHOLD (MG_GRP_DAMAGE_ENABLED + 1)  /* list */
#define MG_GRP_DAMAGE_RESOURCES (MG_GRP_DAMAGE_THRESHOLD + MG_LIST_WORDS) /* list (item, minimum) */
#define MG_GRP_NUM_GOALS (MG_GRP_DAMAGE_RESOURCES + MG_LIST_WORDS)
#define MG_GRP_GOALS (MG_GRP_NUM_GOALS + 1)            /* [MG_MAX_GOALS] token values */
#define MG_GRP_WORDS (MG_GRP_GOALS + MG_MAX_GOALS)

/* ---- CELLTYPES: map cell code (1-based; 0 = empty) -> object config ---- */
#define MG_CELL_KIND 0     /* MG_KIND_* */
#define MG_CELL_TYPE_ID 1  /* reference type_id */
#define MG_CELL_PARAM 2    /* group / asm type / chest type index */
#define MG_CELL_VIBE 3     /* initial vibe */
#define MG_CELL_NUM_TAGS 4
#define MG_CELL_TAGS 5     /* [MG_MAX_TAGS] */
#define MG_CELL_WORDS (MG_CELL_TAGS + MG_MAX_TAGS)
#define MG_KIND_EMPTY 0
#define MG_KIND_WALL 1
#define MG_KIND_AGENT 2
#define MG_KIND_ASSEMBLER 3
#define MG_KIND_CHEST 4

/* ---- ASM_TYPES (assembler_config.hpp:15-36) ---- */
#define MG_ASM_FIRST_PROTOCOL 0
#define MG_ASM_NUM_PROTOCOLS 1
#define MG_ASM_ALLOW_PARTIAL 2
#define MG_ASM_MAX_USES 3
#define MG_ASM_CLIP_IMMUNE 4
#define MG_ASM_START_CLIPPED 5
#define MG_ASM_CHEST_DISTANCE 6
#define MG_ASM_VIBE_DEPENDENT 7  /* current protocol can depend on the neighbourhood */
#define MG_ASM_IDLE_PROTOCOL 8   /* protocol index with no agents around, or -1 */
#define MG_ASM_WORDS 9

/* ---- PROTOCOLS: per assembler type, grouped by key, min_agents descending (assembler.hpp:292-309) ---- */
#define MG_PROTO_KEY_LO 0
#define MG_PROTO_KEY_HI 1
#define MG_PROTO_MIN_AGENTS 2
#define MG_PROTO_NUM_VIBES 3
#define MG_PROTO_VIBES 4                               /* [MG_MAX_PROTO_VIBES] sorted */
#define MG_PROTO_COOLDOWN (MG_PROTO_VIBES + MG_MAX_PROTO_VIBES)
#define MG_PROTO_HAS_OUTPUT (MG_PROTO_COOLDOWN + 1)
#define MG_PROTO_INPUTS (MG_PROTO_HAS_OUTPUT + 1)      /* list, input_resources map order */
#define MG_PROTO_OUTPUTS (MG_PROTO_INPUTS + MG_LIST_WORDS)
#define MG_PROTO_SCALED_INPUTS (MG_PROTO_OUTPUTS + MG_LIST_WORDS)   /* same pairs, order of the rebuilt map */
#define MG_PROTO_SCALED_OUTPUTS (MG_PROTO_SCALED_INPUTS + MG_LIST_WORDS)
#define MG_PROTO_WORDS (MG_PROTO_SCALED_OUTPUTS + MG_LIST_WORDS)

/* ---- CHEST_TYPES (chest_config.hpp:15-31) ---- */
#define MG_CHEST_INVCFG 0
#define MG_CHEST_INITIAL_INV 1                         /* list */
#define MG_CHEST_NUM_VIBES (1 + MG_LIST_WORDS)
#define MG_CHEST_VIBE_IDS (MG_CHEST_NUM_VIBES + 1)     /* [MG_MAX_CHEST_VIBES] */
#define MG_CHEST_VIBE_LISTS (MG_CHEST_VIBE_IDS + MG_MAX_CHEST_VIBES) /* lists (item, delta) */
#define MG_CHEST_WORDS (MG_CHEST_VIBE_LISTS + MG_MAX_CHEST_VIBES * MG_LIST_WORDS)

/* ---- CLIPPER (clipper_config.hpp:11-26) ---- */
#define MG_CLIP_ENABLED 0
#define MG_CLIP_FIRST_PROTOCOL 1
#define MG_CLIP_NUM_PROTOCOLS 2
#define MG_CLIP_LENGTH_SCALE 3
#define MG_CLIP_CUTOFF 4
#define MG_CLIP_PERIOD 5
#define MG_CLIP_WORDS 6

/* ---- INV_FEATURES: [16][MG_MAX_INV_TOKENS] feature id of inv:<r>, inv:<r>:p1.. ---- */
/* ---- PROTO_FEATURES: input[16], output[16] ---- */

/* ---- per-agent stat slots. Resource arrays have stride MG_RES_STRIDE, group arrays MG_MAX_GROUPS ---- */
#define MG_AS_ACTION_SUCCESS 0 /* [3] by action kind */
#define MG_AS_ACTION_FAILED 3  /* [3] */
#define MG_AS_ACTION_FAILED_ANY 6
#define MG_AS_FROZEN_TICKS 7
#define MG_AS_MAX_STEPS_WITHOUT_MOTION 8
#define MG_AS_SWAP 9
#define MG_AS_DIVERSITY 10
#define MG_AS_DAMAGE_ITEMS_LOST 11
#define MG_AS_DAMAGE_TRIGGERED 12
#define MG_AS_INVALID_INDEX 13
#define MG_AS_ATTACK_FRIENDLY_FIRE 14
#define MG_AS_TRANSFER_COUNT 15
#define MG_AS_ATTACK_BLOCKED_BY 16                         /* [MG_MAX_GROUPS] by target config */
#define MG_AS_ATTACK_HIT (16 + MG_MAX_GROUPS)              /* [MG_MAX_GROUPS] */
#define MG_AS_ATTACK_HIT_BY (16 + 2 * MG_MAX_GROUPS)       /* [MG_MAX_GROUPS] */
#define MG_AS_RES_BASE (16 + 3 * MG_MAX_GROUPS)
#define MG_AS_AMOUNT (MG_AS_RES_BASE)                      /* [16] <res>.amount */
#define MG_AS_GAINED (MG_AS_RES_BASE + MG_RES_STRIDE)
#define MG_AS_LOST (MG_AS_RES_BASE + 2 * MG_RES_STRIDE)
#define MG_AS_CHEST_DEPOSITED (MG_AS_RES_BASE + 3 * MG_RES_STRIDE)
#define MG_AS_CHEST_WITHDRAWN (MG_AS_RES_BASE + 4 * MG_RES_STRIDE)
#define MG_AS_DAMAGED (MG_AS_RES_BASE + 5 * MG_RES_STRIDE)
#define MG_AS_TRANSFER_SELF (MG_AS_RES_BASE + 6 * MG_RES_STRIDE)
#define MG_AS_TRANSFER_TO (MG_AS_RES_BASE + 7 * MG_RES_STRIDE) /* [MG_MAX_GROUPS][16] */
#define MG_AS_COUNT (MG_AS_TRANSFER_TO + MG_MAX_GROUPS * MG_RES_STRIDE)

/* ---- per-env game stat slots ---- */
#define MG_GS_TOKENS_WRITTEN 0
#define MG_GS_TOKENS_DROPPED 1
#define MG_GS_TOKENS_FREE_SPACE 2
#define MG_GS_ASM_CREATED 4                            /* [16] assembler.<res>.created */
#define MG_GS_CHEST_DEPOSITED (4 + MG_RES_STRIDE)
#define MG_GS_CHEST_WITHDRAWN (4 + 2 * MG_RES_STRIDE)
#define MG_GS_CHEST_AMOUNT (4 + 3 * MG_RES_STRIDE)
#define MG_GS_OBJECTS (4 + 4 * MG_RES_STRIDE)         /* [MG_MAX_CELLTYPES] objects.<map cell name>, counted at construction (mettagrid_c.cpp:203,224,235) */
#define MG_GS_COUNT (MG_GS_OBJECTS + MG_MAX_CELLTYPES)

/* ---- per-agent record of out-of-range action ids (action.invalid_index.<id>, mettagrid_c.cpp:565-569): the first
 * MG_INVALID_IDS distinct ids an agent sends in an episode with their counts, then a count of all further ones ---- */
#define MG_INVALID_IDS 4
#define MG_INVALID_WORDS (2 * MG_INVALID_IDS + 2) /* (id, count) pairs, count of the rest, pairs in use */

#endif /* MG_CONFIG_H_ */
