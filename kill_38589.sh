ssh test@127.0.0.1 "kill -9 1844"
ssh test@127.0.0.1 "kill -9 1845"
ssh test@127.0.0.1 "kill -9 1846"
ssh test@127.0.0.1 "kill -9 1847"
kill -9 1591
