ssh test@127.0.0.1 "kill -9 5846"
ssh test@127.0.0.1 "kill -9 5847"
kill -9 5044
