ssh test@127.0.0.1 "kill -9 5602"
ssh test@127.0.0.1 "kill -9 5603"
ssh test@127.0.0.1 "kill -9 5604"
kill -9 5044
