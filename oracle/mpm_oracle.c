int orc_mpm_placeholder;
